// eqtl_pair.cuh — the per-pair tail of the dense eQTL scan, shared by the FP64 tensor-core kernel (kernels_eqtl.cu) and the
// tcgen05 int8 kernel (kernels_eqtl_umma.cu): one (snarl, gene) regression from its moments.
#pragma once
#include "device_math.cuh"
#include "eqtl_dense.hpp"
#include "ols.cuh"

namespace stoat {

// what the epilogue needs of EqtlDense, kept in shared memory so that the per-pair function takes one pointer
struct EqConst {
    const double* yc;
    const double* tot;
    const double* tot2;
    const uint8_t* dcode;
    uint64_t s_pad;
    uint32_t n_pad, n_samples;
    double p_threshold, z_threshold;
};

// 2-bit dosage code -> double: 0 -> 0.0, 1 -> 0.5 (0x3FE0...), 2 -> 1.0 (0x3FF0...)
__device__ __forceinline__ double code_to_double(uint32_t code) {
    return __hiloint2double(code ? int(0x3FD00000u + (code << 20)) : 0, 0);
}
// explicit residual sum of squares of one (snarl, gene) pair — linear_regression's own formula (stats_test.cpp:460-463);
// taken when the moment expansion has lost its digits (near-perfect fit).  One thread, rare.
static __device__ __noinline__ double eq_rss_explicit(const double* y, const uint8_t* dcode_s, uint64_t s_pad, uint32_t n_samples,
                                               const EqtlRec* rp, double b0, double b1) {
    const EqtlRec r = *rp;
    double rss = 0.0;
    for (uint32_t i = 0; i < n_samples; ++i) {
        bool unused = false;
        for (uint32_t k = 0; k < r.n_unused; ++k) unused |= (r.unused[k] == i);
        if (unused) continue;
        const uint32_t w = *reinterpret_cast<const uint32_t*>(dcode_s + uint64_t(i / 64) * s_pad * 16 + ((i % 64) / 16) * 4);
        const double d = code_to_double((w >> (2 * (i % 16))) & 3u);
        const double res = y[i] - (b0 + b1 * d);
        rss = fma(res, res, rss);
    }
    return rss;
}

// One (snarl, gene) regression from its moments: the tail of linear_regression (stats_test.cpp:455-489) for X = [1, d].
// Returns (|t| + 1), negated when the pair's p-value is below the threshold; an invalid t (p = 1, :480-482) ranks as t = 0.
static __device__ __noinline__ double eq_pair_stat(const EqConst* k, const EqtlRec* rp, uint64_t s, uint32_t g, double sdy) {
    const EqtlRec& r = *rp;
    const double n_used = double(r.n_used);
    int df_res = int(r.n_used) - 2 + 1;  // stats_test.cpp:467 (sic)
    if (df_res < 1) df_res = 1;
    const double* yrow = k->yc + uint64_t(g) * k->n_pad;
    double sy = k->tot[g], syy = k->tot2[g];
    for (uint32_t u = 0; u < r.n_unused; ++u) {  // samples on no path are not rows of the table
        const double y = yrow[r.unused[u]];
        sy -= y;
        syy = fma(-y, y, syy);
    }
    const double b0 = r.inv00 * sy + r.inv01 * sdy, b1 = r.inv01 * sy + r.inv11 * sdy;
    const double bxy = b0 * sy + b1 * sdy;
    const double bxxb = b0 * (n_used * b0 + r.sd * b1) + b1 * (r.sd * b0 + r.sdd * b1);
    double rss = syy - 2.0 * bxy + bxxb;
    const double tss = syy - sy * sy / n_used;
    if (!(rss > 1e-6 * tss)) rss = eq_rss_explicit(yrow, k->dcode + s * 16, k->s_pad, k->n_samples, rp, b0, b1);
    const double se = sqrt(r.inv11 * (rss / double(df_res)));
    const double t = b1 / se;
    const bool fin = !(isnan(t) || isinf(t));
    const double key = fin ? fabs(t) : 0.0;
    const bool below = fin ? (key > k->z_threshold && 2.0 * student_t_upper_tail(double(df_res), key) < k->p_threshold)
                           : (1.0 < k->p_threshold);
    return below ? -(key + 1.0) : key + 1.0;
}

}  // namespace stoat
