// device_math.cuh — FP64 device routines for the statistics epilogues (K2 binary_stats,
// K4 ols_solve).  Everything the reference takes from Boost.Math is evaluated here in double:
//   chi-squared upper tail  (src/stats_test.cpp:205-210, 246-256)
//   Student-t upper tail    (src/stats_test.cpp:476-484)
// and the PLINK-style Fisher recurrence (src/stats_test.cpp:264-360) is restated with explicit
// round-to-nearest intrinsics so that ptxas can never contract it into FMAs: the reference's
// x86-64 build (-O3, no -march: CMakeLists.txt:34-35) performs plain IEEE mul/div/add, and the
// recurrence's tie test (cur_prob < kExactTestBias, tprob <= preaddp) makes the result depend on
// every rounding.
#pragma once
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cuda_runtime.h>

namespace stoat {

// codes of include/stoat_gpu.h
enum : uint8_t { kPValue = 0, kPNa = 1, kPZero = 2, kPOne = 3, kPDblMax = 4, kPUndef = 5, kPOverLimit = 6, kPDeferred = 7 };

struct PVal {
    double v;
    uint8_t code;
};

// boost::math::cdf(normal_distribution<>(0, 1), z) = erfc(-z / sqrt 2) / 2 (src/stats_test.cpp:22-25); the logistic
// regression's Wald test then forms 2 (1 - cdf), a multiple of 2^-52 (:151-152)
__device__ inline double normal_cdf_like_reference(double z) { return 0.5 * erfc(-z / 1.4142135623730951); }

// ------------------------------------------------------------------------------------------------
// Regularized upper incomplete gamma Q(a, x) for a = df/2, df a positive integer — finite sums:
//   Q(n, x)     = e^-x * sum_{k<n} x^k / k!
//   Q(n+1/2, x) = erfc(sqrt x) + e^-x * sum_{j<n} x^(j+1/2) / Gamma(j+3/2)
// (all terms positive: no cancellation).  These are also the forms Boost's gamma_q takes for
// integer / half-integer a < 30.
// ------------------------------------------------------------------------------------------------
__device__ inline double gamma_q_half_integer_df(unsigned df, double x) {
    if (!(x > 0.0)) return 1.0;
    double sum, base;
    unsigned terms;
    if (df & 1u) {  // a = n + 1/2, n = df / 2
        base = erfc(sqrt(x));
        terms = df >> 1;
        if (terms == 0) return base;
        // first term x^(1/2) / Gamma(3/2) = 2 sqrt(x / pi)
        double t = 2.0 * sqrt(x * 0.31830988618379067154);
        sum = t;
        for (unsigned j = 1; j < terms; ++j) {
            t *= x / (double(j) + 0.5);
            sum += t;
        }
    } else {  // a = n = df / 2 >= 1
        base = 0.0;
        terms = df >> 1;
        double t = 1.0;
        sum = 1.0;
        for (unsigned k = 1; k < terms; ++k) {
            t *= x / double(k);
            sum += t;
        }
    }
    const double scaled = (x < 700.0) ? exp(-x) * sum : exp(log(sum) - x);
    return base + scaled;
}

// The reference's p-value: `1.0 - cdf` evaluated in double when stat <= 85 (so p is quantised to
// the double grid below 1.0 — golden rows print "0" for p ~ 1e-19), the true upper tail (50-digit
// arithmetic in the reference) when stat > 85.
__device__ inline double chi2_pvalue_like_reference(double stat, unsigned df) {
    const double x = 0.5 * stat;
    if (df == 1u && x <= 0.2) {
        // small argument: the lower tail is the directly computed quantity
        return 1.0 - erf(sqrt(x));
    }
    const double q = gamma_q_half_integer_df(df, x);
    if (stat > 85.0) return q;
    const double cdf = 1.0 - q;  // what boost::math::cdf returns (gamma_p = 1 - gamma_q)
    return 1.0 - cdf;
}

// ------------------------------------------------------------------------------------------------
// Student-t upper tail  P(T > t), t >= 0, df >= 1:  0.5 * I_x(df/2, 1/2), x = df / (df + t^2),
// by the continued fraction of the incomplete beta (modified Lentz) on whichever side converges.
// ------------------------------------------------------------------------------------------------
__device__ inline double ibeta_cf(double a, double b, double x) {
    const double tiny = 1e-300;
    const double qab = a + b, qap = a + 1.0, qam = a - 1.0;
    double c = 1.0;
    double d = 1.0 - qab * x / qap;
    if (fabs(d) < tiny) d = tiny;
    d = 1.0 / d;
    double h = d;
    for (int m = 1; m <= 2000; ++m) {
        const double m2 = 2.0 * m;
        double aa = m * (b - m) * x / ((qam + m2) * (a + m2));
        d = 1.0 + aa * d;
        if (fabs(d) < tiny) d = tiny;
        c = 1.0 + aa / c;
        if (fabs(c) < tiny) c = tiny;
        d = 1.0 / d;
        h *= d * c;
        aa = -(a + m) * (qab + m) * x / ((a + m2) * (qap + m2));
        d = 1.0 + aa * d;
        if (fabs(d) < tiny) d = tiny;
        c = 1.0 + aa / c;
        if (fabs(c) < tiny) c = tiny;
        d = 1.0 / d;
        const double delta = d * c;
        h *= delta;
        if (fabs(delta - 1.0) < 2e-16) break;
    }
    return h;
}

// The same continued fraction one iteration at a time (K4b runs it on persistent lanes that pick up the next snarl as
// soon as theirs has converged: the iteration count varies from ~5 to hundreds with x).  init + step() until it returns
// true reproduce ibeta_cf() operation for operation.
struct IbetaCf {
    double a, b, x, qab, qap, qam, c, d, h;
    int m;
    __device__ void init(double a_, double b_, double x_) {
        const double tiny = 1e-300;
        a = a_; b = b_; x = x_;
        qab = a + b; qap = a + 1.0; qam = a - 1.0;
        c = 1.0;
        d = 1.0 - qab * x / qap;
        if (fabs(d) < tiny) d = tiny;
        d = 1.0 / d;
        h = d;
        m = 1;
    }
    // resume after init(): d0 = the d (= h) init() computes (K4 does that division with all lanes busy)
    __device__ void init_from(double a_, double b_, double x_, double d0) {
        a = a_; b = b_; x = x_;
        qab = a + b; qap = a + 1.0; qam = a - 1.0;
        c = 1.0;
        d = d0;
        h = d0;
        m = 1;
    }
    __device__ bool step() {  // true: converged (or out of iterations), h is the value
        const double tiny = 1e-300;
        const double m2 = 2.0 * m;
        double aa = m * (b - m) * x / ((qam + m2) * (a + m2));
        d = 1.0 + aa * d;
        if (fabs(d) < tiny) d = tiny;
        c = 1.0 + aa / c;
        if (fabs(c) < tiny) c = tiny;
        d = 1.0 / d;
        h *= d * c;
        aa = -(a + m) * (qab + m) * x / ((a + m2) * (qap + m2));
        d = 1.0 + aa * d;
        if (fabs(d) < tiny) d = tiny;
        c = 1.0 + aa / c;
        if (fabs(c) < tiny) c = tiny;
        d = 1.0 / d;
        const double delta = d * c;
        h *= delta;
        ++m;
        return fabs(delta - 1.0) < 2e-16 || m > 2000;
    }
};

// ln B(a, 1/2) = lgamma(a) + lgamma(1/2) - lgamma(a + 1/2); asymptotic difference for large a so
// that two ~1e5-sized lgamma values are never subtracted.
__device__ inline double log_beta_half(double a) {
    const double half_log_pi = 0.57236494292470008707;  // ln Gamma(1/2)
    if (a < 40.0) return lgamma(a) + half_log_pi - lgamma(a + 0.5);
    // ln Gamma(a) - ln Gamma(a + 1/2) = -1/2 ln a + 1/(8a) - 1/(192 a^3) + 1/(640 a^5) - 17/(14336 a^7) + ...
    const double r = 1.0 / a, r2 = r * r;
    const double series = r * (0.125 + r2 * (-1.0 / 192.0 + r2 * (1.0 / 640.0 + r2 * (-17.0 / 14336.0))));
    return half_log_pi - 0.5 * log(a) + series;
}

__device__ inline double student_t_upper_tail(double df, double t) {
    if (isinf(t)) return 0.0;
    const double t2 = t * t;
    if (t2 == 0.0) return 0.5;
    const double a = 0.5 * df, b = 0.5;
    const double z = t2 / (df + t2);  // 1 - x
    // x^a (1-x)^b / B(a,b) with x^a = (1 + t^2/df)^-a evaluated through log1p
    const double log_front = -a * log1p(t2 / df) + b * log(z) - log_beta_half(a);
    const double front = exp(log_front);
    if (z > 1.5 / (a + 2.5)) {
        const double x = df / (df + t2);
        return 0.5 * front * ibeta_cf(a, b, x) / a;
    }
    return 0.5 * (1.0 - front * ibeta_cf(b, a, z) / b);
}

// ------------------------------------------------------------------------------------------------
// FisherKhi2::chi2_2x2 — src/stats_test.cpp:181-211.  Statistic in the reference's operation
// order with explicit _rn intrinsics.
// ------------------------------------------------------------------------------------------------
__device__ inline PVal chi2_2x2(uint64_t a, uint64_t b, uint64_t c, uint64_t d) {
    const long long row1 = (long long)(a + b), row2 = (long long)(c + d);
    const long long col1 = (long long)(a + c), col2 = (long long)(b + d);
    const long long total = row1 + row2;
    if (row1 == 0 || row2 == 0 || col1 == 0 || col2 == 0) return {0.0, kPNa};
    const double dt = double(total);
    const double ea = __ddiv_rn(__dmul_rn(double(row1), double(col1)), dt);
    const double eb = __ddiv_rn(__dmul_rn(double(row1), double(col2)), dt);
    const double ec = __ddiv_rn(__dmul_rn(double(col1), double(row2)), dt);
    const double ed = __ddiv_rn(__dmul_rn(double(col2), double(row2)), dt);
    if (ea == 0 || eb == 0 || ec == 0 || ed == 0) return {DBL_MAX, kPDblMax};
    double stat = 0.0, t;
    t = __dsub_rn(double(a), ea); stat = __dadd_rn(stat, __ddiv_rn(__dmul_rn(t, t), ea));
    t = __dsub_rn(double(b), eb); stat = __dadd_rn(stat, __ddiv_rn(__dmul_rn(t, t), eb));
    t = __dsub_rn(double(c), ec); stat = __dadd_rn(stat, __ddiv_rn(__dmul_rn(t, t), ec));
    t = __dsub_rn(double(d), ed); stat = __dadd_rn(stat, __ddiv_rn(__dmul_rn(t, t), ed));
    return {chi2_pvalue_like_reference(stat, 1u), kPValue};
}

// one column's contribution to the 2xN statistic — src/stats_test.cpp:238-244
__device__ inline double chi2_2xN_column(double stat, uint64_t g0, uint64_t g1, uint64_t row0, uint64_t row1,
                                         uint64_t total) {
    const double ct = double(g0 + g1), dt = double(total);
    const double e0 = __ddiv_rn(__dmul_rn(double(row0), ct), dt);
    const double e1 = __ddiv_rn(__dmul_rn(double(row1), ct), dt);
    const double d0 = __dsub_rn(double(g0), e0);
    const double d1 = __dsub_rn(double(g1), e1);
    stat = __dadd_rn(stat, __ddiv_rn(__dmul_rn(d0, d0), e0));
    stat = __dadd_rn(stat, __ddiv_rn(__dmul_rn(d1, d1), e1));
    return stat;
}

// ------------------------------------------------------------------------------------------------
// FisherKhi2::fastFishersExactTest — src/stats_test.cpp:264-360
// ------------------------------------------------------------------------------------------------
// The reference runs three loops (towards the centre, far tail, near tail).  All three have the same body
//   d1 += 1; d2 += 1; cur_prob *= (n1 * n2) / (d1 * d2); n1 -= 1; n2 -= 1;
// with (n, d) = (cur12, cur21 | cur11, cur22) in the first two and (cur11, cur22 | cur12, cur21) in the third.
// They are fused into ONE loop with a per-thread phase so that the lanes of a warp — each on its own table, in
// whatever phase — share the expensive instructions (3 DMUL + 1 DDIV) instead of serialising three loops.
// Every floating-point operation and its order are those of the reference.
__device__ inline PVal fisher_2x2(uint64_t m11, uint64_t m12, uint64_t m21, uint64_t m22) {
    if ((m11 | m12) == 0 || (m21 | m22) == 0 || (m11 | m21) == 0 || (m12 | m22) == 0) return {0.0, kPNa};
    constexpr double kEps2 = 9.094947017729282e-13;  // stats_test.hpp:59
    constexpr double kBias = 1.0339757656912846e-25;  // stats_test.hpp:60
    constexpr double kStart = (1 - kEps2) * kBias;    // folded in IEEE double, as on the host
    uint64_t u;
    if (m12 > m21) { u = m12; m12 = m21; m21 = u; }  // :280-284
    if (m11 > m22) { u = m11; m11 = m22; m22 = u; }  // :286-290
    if (m11 * m22 > m12 * m21) {                      // :292-299
        u = m11; m11 = m12; m12 = u;
        u = m21; m21 = m22; m22 = u;
    }
    double tprob = kStart, cur_prob = kStart, cprob = 0.0;
    double n1 = double(m12), n2 = double(m21), d1 = double(m11), d2 = double(m22);
    int phase = 1;
    PVal res{0.0, kPNa};
    for (;;) {
        // ---- loop conditions / transitions evaluated before a body execution ----------------------
        if (phase == 1 && !(n1 > 0.5)) {  // first loop ran out (:306)
            if (cprob == 0) { res = {1.0, kPOne}; break; }  // :322-324
            phase = 2;
        }
        if (phase == 2 && !(n1 > 0.5)) phase = 4;  // second loop's condition (:326)
        if (phase == 4) {                          // set up the third loop (:339-344)
            if (m11 == 0) { res = {__ddiv_rn(tprob, __dadd_rn(cprob, tprob)), kPValue}; break; }  // :359
            n1 = double(m11); n2 = double(m22); d1 = double(m12); d2 = double(m21);
            cur_prob = kStart;
            phase = 3;
        }
        // ---- common body ----------------------------------------------------------------------------
        d1 += 1; d2 += 1;  // exact: small integers
        cur_prob = __dmul_rn(cur_prob, __ddiv_rn(__dmul_rn(n1, n2), __dmul_rn(d1, d2)));
        n1 -= 1; n2 -= 1;
        // ---- per-phase bookkeeping ---------------------------------------------------------------------
        if (phase == 1) {
            if (cur_prob > DBL_MAX) { res = {0.0, kPZero}; break; }  // :312-314
            if (cur_prob < kBias) {                                   // :315-318
                tprob = __dadd_rn(tprob, cur_prob);
                if (cprob == 0) { res = {1.0, kPOne}; break; }
                phase = 2;
            } else {
                cprob = __dadd_rn(cprob, cur_prob);
            }
        } else {
            const double preaddp = tprob;
            tprob = __dadd_rn(tprob, cur_prob);
            if (phase == 2) {
                if (tprob <= preaddp) phase = 4;  // :334-336
            } else {
                if (tprob <= preaddp) { res = {__ddiv_rn(preaddp, __dadd_rn(cprob, preaddp)), kPValue}; break; }  // :352-354
                if (!(n1 > 0.5)) { res = {__ddiv_rn(tprob, __dadd_rn(cprob, tprob)), kPValue}; break; }           // :356, :359
            }
        }
    }
    return res;
}

}  // namespace stoat
