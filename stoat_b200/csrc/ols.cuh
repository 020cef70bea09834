// ols.cuh — small dense FP64 linear algebra + the regression epilogue shared by K4 (ols_solve) and
// the general quantitative path.  Follows LinearRegression::computeXtXinverse / linear_regression
// (src/stats_test.cpp:398-506): (X^T X)^-1 by LDLT with diagonal pivoting (what Eigen::LDLT does),
// falling back to the pseudo-inverse (singular values below 1e-6 dropped) when any |D_i| < 1e-10.
#pragma once
#include "device_math.cuh"

namespace stoat {

// symmetric matrix helpers, column-major a[i + j*N]
template <int N>
struct SmallMat {
    double a[N * N];
    __device__ double& operator()(int i, int j) { return a[i + j * N]; }
    __device__ double operator()(int i, int j) const { return a[i + j * N]; }
};

// In-place pivoted LDL^T of the leading n x n block (lower triangle is referenced and overwritten
// by L; D on the diagonal).  tr[k] = index swapped with k at step k.
template <int N>
__device__ inline void ldlt_factor(SmallMat<N>& m, int n, int* tr) {
    for (int k = 0; k < n; ++k) {
        int piv = k;
        double best = fabs(m(k, k));
        for (int i = k + 1; i < n; ++i) {
            const double v = fabs(m(i, i));
            if (v > best) {
                best = v;
                piv = i;
            }
        }
        tr[k] = piv;
        if (piv != k) {
            for (int j = 0; j < k; ++j) {
                const double t = m(k, j);
                m(k, j) = m(piv, j);
                m(piv, j) = t;
            }
            for (int i = piv + 1; i < n; ++i) {
                const double t = m(i, k);
                m(i, k) = m(i, piv);
                m(i, piv) = t;
            }
            {
                const double t = m(k, k);
                m(k, k) = m(piv, piv);
                m(piv, piv) = t;
            }
            for (int i = k + 1; i < piv; ++i) {
                const double t = m(i, k);
                m(i, k) = m(piv, i);
                m(piv, i) = t;
            }
        }
        double dkk = m(k, k);
        for (int j = 0; j < k; ++j) dkk -= m(k, j) * (m(j, j) * m(k, j));
        m(k, k) = dkk;
        for (int i = k + 1; i < n; ++i) {
            double s = 0.0;
            for (int j = 0; j < k; ++j) s += m(i, j) * (m(j, j) * m(k, j));
            m(i, k) -= s;
        }
        if (fabs(dkk) > 0.0)
            for (int i = k + 1; i < n; ++i) m(i, k) /= dkk;
    }
}

template <int N>
__device__ inline void ldlt_solve(const SmallMat<N>& m, int n, const int* tr, double* b) {
    for (int k = 0; k < n; ++k) {
        const double t = b[k];
        b[k] = b[tr[k]];
        b[tr[k]] = t;
    }
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < i; ++j) b[i] -= m(i, j) * b[j];
    for (int i = 0; i < n; ++i) b[i] = (fabs(m(i, i)) > DBL_MIN) ? b[i] / m(i, i) : 0.0;
    for (int i = n - 1; i >= 0; --i)
        for (int j = i + 1; j < n; ++j) b[i] -= m(j, i) * b[j];
    for (int k = n - 1; k >= 0; --k) {
        const double t = b[k];
        b[k] = b[tr[k]];
        b[tr[k]] = t;
    }
}

// pseudo-inverse of a symmetric matrix through cyclic Jacobi: sum over |lambda| > tol of v v^T / lambda
template <int N>
__device__ inline void pinv_symmetric(const SmallMat<N>& A, int n, double tol, SmallMat<N>& out) {
    SmallMat<N> S = A, V;
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < n; ++j) V(i, j) = (i == j) ? 1.0 : 0.0;
    for (int sweep = 0; sweep < 60; ++sweep) {
        double off = 0.0;
        for (int p = 0; p < n; ++p)
            for (int q = p + 1; q < n; ++q) off += S(p, q) * S(p, q);
        if (off == 0.0) break;
        for (int p = 0; p < n; ++p)
            for (int q = p + 1; q < n; ++q) {
                if (S(p, q) == 0.0) continue;
                const double theta = (S(q, q) - S(p, p)) / (2.0 * S(p, q));
                const double t = (theta >= 0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
                const double c = 1.0 / sqrt(t * t + 1.0), s = t * c;
                for (int k = 0; k < n; ++k) {
                    const double a = S(k, p), b = S(k, q);
                    S(k, p) = c * a - s * b;
                    S(k, q) = s * a + c * b;
                }
                for (int k = 0; k < n; ++k) {
                    const double a = S(p, k), b = S(q, k);
                    S(p, k) = c * a - s * b;
                    S(q, k) = s * a + c * b;
                }
                for (int k = 0; k < n; ++k) {
                    const double a = V(k, p), b = V(k, q);
                    V(k, p) = c * a - s * b;
                    V(k, q) = s * a + c * b;
                }
            }
    }
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < n; ++j) out(i, j) = 0.0;
    for (int e = 0; e < n; ++e) {
        const double lam = S(e, e);
        if (fabs(lam) > tol)
            for (int i = 0; i < n; ++i)
                for (int j = 0; j < n; ++j) out(i, j) += V(i, e) * V(j, e) / lam;
    }
}

// computeXtXinverse (src/stats_test.cpp:398-421): XtX full symmetric n x n in, inverse out.
// Returns true when the rank-deficient (pseudo-inverse) branch was taken.
template <int N>
__device__ inline bool xtx_inverse(const SmallMat<N>& XtX, int n, SmallMat<N>& inv) {
    SmallMat<N> f = XtX;
    int tr[N];
    ldlt_factor<N>(f, n, tr);
    bool deficient = false;
    for (int i = 0; i < n; ++i)
        if (fabs(f(i, i)) < 1e-10) deficient = true;
    if (deficient) {
        pinv_symmetric<N>(XtX, n, 1e-6, inv);
        return true;
    }
    double e[N];
    for (int j = 0; j < n; ++j) {
        for (int i = 0; i < n; ++i) e[i] = (i == j) ? 1.0 : 0.0;
        ldlt_solve<N>(f, n, tr, e);
        for (int i = 0; i < n; ++i) inv(i, j) = e[i];
    }
    return false;
}

// Diagonal of the inverse of a general n x n matrix by row-pivoted LU and n unit-vector solves: what Eigen's
// MatrixXd::inverse() does for a dynamic-size matrix (PartialPivLU), used by the logistic regression for the covariance
// (src/stats_test.cpp:138-139).
template <int N>
__device__ inline void lu_inverse_diag(const SmallMat<N>& A, int n, double* diag) {
    SmallMat<N> lu = A;
    int perm[N];
    for (int i = 0; i < n; ++i) perm[i] = i;
    for (int k = 0; k < n; ++k) {
        int piv = k;
        double best = fabs(lu(k, k));
        for (int i = k + 1; i < n; ++i) {
            const double v = fabs(lu(i, k));
            if (v > best) {
                best = v;
                piv = i;
            }
        }
        if (piv != k) {
            for (int j = 0; j < n; ++j) {
                const double t = lu(k, j);
                lu(k, j) = lu(piv, j);
                lu(piv, j) = t;
            }
            const int t = perm[k];
            perm[k] = perm[piv];
            perm[piv] = t;
        }
        const double d = lu(k, k);
        if (d != 0.0)
            for (int i = k + 1; i < n; ++i) lu(i, k) /= d;
        for (int j = k + 1; j < n; ++j)
            for (int i = k + 1; i < n; ++i) lu(i, j) -= lu(i, k) * lu(k, j);
    }
    double b[N];
    for (int c = 0; c < n; ++c) {
        for (int i = 0; i < n; ++i) b[i] = (perm[i] == c) ? 1.0 : 0.0;
        for (int i = 0; i < n; ++i)
            for (int j = 0; j < i; ++j) b[i] -= lu(i, j) * b[j];
        for (int i = n - 1; i >= c; --i) {  // rows below c of column c are not needed for the diagonal
            for (int j = i + 1; j < n; ++j) b[i] -= lu(i, j) * b[j];
            b[i] /= lu(i, i);
        }
        diag[c] = b[c];
    }
}

struct OlsOut {
    double p, r2, beta, se;
};

// the tail of linear_regression (src/stats_test.cpp:460-489) once beta, (XtX)^-1[1][1], rss and tss
// are known
__device__ inline OlsOut ols_statistics(double beta1, double inv11, double rss, double tss, uint32_t n_rows, int p) {
    OlsOut o;
    o.r2 = 1.0 - (rss / tss);
    int df_res = int(n_rows) - p + 1;  // :467 (sic)
    if (df_res < 1) df_res = 1;        // :468
    const double mse = rss / double(df_res);
    o.se = sqrt(inv11 * mse);
    o.beta = beta1;
    const double t = beta1 / o.se;
    if (isnan(t) || isinf(t))
        o.p = 1.0;  // :480-482
    else
        o.p = 2.0 * student_t_upper_tail(double(df_res), fabs(t));  // :484
    return o;
}

}  // namespace stoat
