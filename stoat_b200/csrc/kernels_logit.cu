// kernels_logit.cu — the binary trait with covariates: BinaryCovarSnarlAnalyzer::analyze_and_write_snarl
// (src/snarl_analyzer.cpp:392-428) -> LogisticRegression::logistic_regression (src/stats_test.cpp:49-176).  SURVEY §8f-2.
//
// The reference builds the quantitative dosage table against the BINARY phenotype, applies the same column surgery and
// filters as the quantitative analyzer and then fits a ridge-penalised logistic model by IRLS over the n_used samples.
// As written the covariates never enter the design (the block at stats_test.cpp:59-62 declares new locals that shadow
// num_covariates / num_features), so X = [1, dosage columns].
//
// B200 design: a two-path snarl (93 % of a cohort) has ONE dosage column with at most five distinct values
// (0, 1/3, 1/2, 2/3, 1), so the whole regression is a function of TEN integers — samples per (dosage class, phenotype):
//   L1 logit_class_counts   warp per snarl, 128-bit row loads, the class masks of dosage_masks.cuh, popcounts of
//                           (class) and (class & case): HBM bound like K1, exact integers
//   L2 logit_irls           thread per snarl: filters (snarl_analyzer.cpp:542-579) from the counts, IRLS over the <= 5
//                           weighted points (2 x 2 pivoted LDLT per step, as Eigen::LDLT), covariance by LU, Wald p
// Any other shape (three or more kept columns -> Holm over the columns, identical columns, > 8 checked edges, filter
// comparisons of non-dyadic sums within rounding of a threshold) goes to the general kernel's logistic mode
// (quant_general.cuh, per-sample sweeps), which also serves `use_covariates` — the design the reference evidently meant.
#include <algorithm>

#include "ctx.hpp"
#include "dosage_masks.cuh"
#include "ols.cuh"

namespace stoat {

namespace {

constexpr uint32_t kMissingL = 0xFFFFFFFFu;
constexpr int kLogitMaxEdges = 8;       // checked edges per path on the class-count path
constexpr uint32_t kLogitGeneral = 0xFFFFFFFFu;  // LogitCounts::al0 of a snarl handed to the general kernel

__device__ __forceinline__ uint4 ldg_nc_v4(const uint8_t* p) {
    uint4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(p));
    return v;
}

// three paths of at most two checked edges each: logit_three_path_kernel (L3) takes the snarl
__device__ __forceinline__ bool logit_three_shape(const BatchDev& b, uint64_t p0, uint64_t p1) {
    if (p1 - p0 != 3) return false;
    for (int j = 0; j < 3; ++j)
        if (b.path_edge_off[p0 + j + 1] - b.path_edge_off[p0 + j] > 2) return false;
    return true;
}

struct Acc12 {
    uint32_t al0 = 0, al1 = 0, used = 0, full0 = 0, half = 0, third = 0, tthird = 0;
    uint32_t c_used = 0, c_full0 = 0, c_half = 0, c_third = 0, c_tthird = 0;
};

__device__ __forceinline__ void logit_word(Acc12& a, uint32_t v0, uint32_t v1, uint32_t cm) {
    a.al0 += __popc(v0);
    a.al1 += __popc(v1);
    const LeanMasks m = lean_masks(v0, v1);
    a.used += __popc(m.used);
    a.full0 += __popc(m.full0);
    a.half += __popc(m.half);
    a.third += __popc(m.third);
    a.tthird += __popc(m.tthird);
    a.c_used += __popc(m.used & cm);
    a.c_full0 += __popc(m.full0 & cm);
    a.c_half += __popc(m.half & cm);
    a.c_third += __popc(m.third & cm);
    a.c_tthird += __popc(m.tthird & cm);
}

}  // namespace

// ---- L1: class x phenotype counts of the two-path snarls -------------------------------------------------------------
__global__ void __launch_bounds__(256) logit_class_counts_kernel(const uint8_t* __restrict__ rows, uint32_t pitch,
                                                                 uint32_t n_hap, const uint8_t* __restrict__ case_mask,
                                                                 const BatchDev b, LogitCounts* __restrict__ out,
                                                                 uint32_t* __restrict__ general_list,
                                                                 uint32_t* __restrict__ n_general) {
    const uint32_t lane = threadIdx.x & 31;
    const uint64_t warp = (uint64_t(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const uint64_t n_warps = (uint64_t(gridDim.x) * blockDim.x) >> 5;
    const uint32_t chunks = pitch / 16;
    for (uint64_t s = warp; s < b.n_snarls; s += n_warps) {
        const uint64_t p0 = b.snarl_path_off[s], p1 = b.snarl_path_off[s + 1];
        bool general = (p1 - p0) != 2;
        uint64_t e0 = 0, e1 = 0, e2 = 0;
        if (!general) {
            e0 = b.path_edge_off[p0];
            e1 = b.path_edge_off[p0 + 1];
            e2 = b.path_edge_off[p0 + 2];
            general = (e1 - e0) > uint64_t(kLogitMaxEdges) || (e2 - e1) > uint64_t(kLogitMaxEdges);
        }
        if (general) {
            if (lane == 0) {
                out[s].al0 = kLogitGeneral;
                if (!logit_three_shape(b, p0, p1)) general_list[atomicAdd(n_general, 1u)] = uint32_t(s);  // else L3 has it
            }
            continue;
        }
        const uint32_t r0 = uint32_t(e1 - e0), r1 = uint32_t(e2 - e1);
        // lane j < r0 holds row j of path 0, lane 8 + j row j of path 1
        uint32_t my_row = 0;
        if (lane < r0) my_row = b.edge_row[e0 + lane];
        if (lane >= 8 && lane - 8 < r1) my_row = b.edge_row[e1 + (lane - 8)];
        // a missing edge empties the path (identify_path returns nothing, src/snarl_analyzer.cpp:321-337)
        const uint32_t miss = __ballot_sync(0xffffffffu, my_row == kMissingL && (lane < r0 || (lane >= 8 && lane - 8 < r1)));
        const bool none0 = (miss & 0xFFu) != 0, none1 = (miss & 0xFF00u) != 0;
        // every lane keeps all row indices (the chunk loop below is not warp-uniform in its last round)
        uint32_t rw0[kLogitMaxEdges], rw1[kLogitMaxEdges];
#pragma unroll
        for (int j = 0; j < kLogitMaxEdges; ++j) {
            rw0[j] = __shfl_sync(0xffffffffu, my_row, j);
            rw1[j] = __shfl_sync(0xffffffffu, my_row, 8 + j);
        }
        Acc12 a;
        for (uint32_t c = lane; c < chunks; c += 32) {
            const uint4 cm4 = *reinterpret_cast<const uint4*>(case_mask + size_t(c) * 16);
            const uint4 all = make_uint4(valid_word(4 * c, n_hap), valid_word(4 * c + 1, n_hap), valid_word(4 * c + 2, n_hap),
                                         valid_word(4 * c + 3, n_hap));  // no checked edge: every haplotype
            uint4 v0 = none0 ? make_uint4(0, 0, 0, 0) : all;
            uint4 v1 = none1 ? make_uint4(0, 0, 0, 0) : all;
            if (!none0) {
#pragma unroll
                for (int j = 0; j < kLogitMaxEdges; ++j)
                    if (uint32_t(j) < r0) {
                        const uint4 w = ldg_nc_v4(rows + size_t(rw0[j]) * pitch + size_t(c) * 16);
                        v0.x &= w.x; v0.y &= w.y; v0.z &= w.z; v0.w &= w.w;
                    }
            }
            if (!none1) {
#pragma unroll
                for (int j = 0; j < kLogitMaxEdges; ++j)
                    if (uint32_t(j) < r1) {
                        const uint4 w = ldg_nc_v4(rows + size_t(rw1[j]) * pitch + size_t(c) * 16);
                        v1.x &= w.x; v1.y &= w.y; v1.z &= w.z; v1.w &= w.w;
                    }
            }
            logit_word(a, v0.x, v1.x, cm4.x);
            logit_word(a, v0.y, v1.y, cm4.y);
            logit_word(a, v0.z, v1.z, cm4.z);
            logit_word(a, v0.w, v1.w, cm4.w);
        }
        __syncwarp();
        LogitCounts r;
        if (n_hap < 65536u) {  // every total fits 16 bits: two counts per reduction
            const uint32_t q0 = __reduce_add_sync(0xffffffffu, a.al0 | a.al1 << 16);
            const uint32_t q1 = __reduce_add_sync(0xffffffffu, a.used | a.c_used << 16);
            const uint32_t q2 = __reduce_add_sync(0xffffffffu, a.full0 | a.c_full0 << 16);
            const uint32_t q3 = __reduce_add_sync(0xffffffffu, a.half | a.c_half << 16);
            const uint32_t q4 = __reduce_add_sync(0xffffffffu, a.third | a.c_third << 16);
            const uint32_t q5 = __reduce_add_sync(0xffffffffu, a.tthird | a.c_tthird << 16);
            r.al0 = q0 & 0xFFFFu; r.al1 = q0 >> 16;
            r.used = q1 & 0xFFFFu; r.c_used = q1 >> 16;
            r.full0 = q2 & 0xFFFFu; r.c_full0 = q2 >> 16;
            r.half = q3 & 0xFFFFu; r.c_half = q3 >> 16;
            r.third = q4 & 0xFFFFu; r.c_third = q4 >> 16;
            r.tthird = q5 & 0xFFFFu; r.c_tthird = q5 >> 16;
        } else {
            r.al0 = __reduce_add_sync(0xffffffffu, a.al0); r.al1 = __reduce_add_sync(0xffffffffu, a.al1);
            r.used = __reduce_add_sync(0xffffffffu, a.used); r.c_used = __reduce_add_sync(0xffffffffu, a.c_used);
            r.full0 = __reduce_add_sync(0xffffffffu, a.full0); r.c_full0 = __reduce_add_sync(0xffffffffu, a.c_full0);
            r.half = __reduce_add_sync(0xffffffffu, a.half); r.c_half = __reduce_add_sync(0xffffffffu, a.c_half);
            r.third = __reduce_add_sync(0xffffffffu, a.third); r.c_third = __reduce_add_sync(0xffffffffu, a.c_third);
            r.tthird = __reduce_add_sync(0xffffffffu, a.tthird); r.c_tthird = __reduce_add_sync(0xffffffffu, a.c_tthird);
        }
        if (lane == 0) out[s] = r;
    }
}

// ---- L2: filters + IRLS on the class counts ---------------------------------------------------------------------------
namespace {

struct Logit2 {
    double p, beta, se;
    uint8_t code;
};

// LogisticRegression::logistic_regression (src/stats_test.cpp:49-176) for X = [1, d] whose rows take the values d[c] with
// n0[c] controls and n1[c] cases.  Same operations per sample as the reference (sigmoid, clamped weight, rows scaled by
// sqrt(w) before the products), the sums over samples taken as count x term.
__device__ Logit2 logit_irls_classes(const double (&d)[5], const uint32_t (&n0)[5], const uint32_t (&n1)[5]) {
    constexpr double kTol = 1e-6, kL2 = 1e-4, kEps = 1e-8;  // src/stats_test.hpp:109-112
    constexpr int kMaxIter = 100;
    const double nan = __longlong_as_double(0x7ff8000000000000LL);
    double b0 = 0.0, b1 = 0.0, o0 = 0.0, o1 = 0.0;
    bool converged = false;
    for (int pass = 0; pass <= kMaxIter; ++pass) {
        double h00 = 0.0, h01 = 0.0, h11 = 0.0, g0 = 0.0, g1 = 0.0;
#pragma unroll
        for (int c = 0; c < 5; ++c) {
            const uint32_t nt = n0[c] + n1[c];
            if (!nt) continue;
            const double z = fma(d[c], b1, b0);
            const double pr = 1.0 / (1.0 + exp(-z));                  // :27-29
            const double w = fmax(kEps, fmin(1.0, pr * (1.0 - pr)));   // :93
            const double sw = sqrt(w), x1 = d[c] * sw;                 // :97-98
            const double cnt = double(nt);
            h00 = fma(cnt, sw * sw, h00);
            h01 = fma(cnt, x1 * sw, h01);
            h11 = fma(cnt, x1 * x1, h11);
            const double r = fma(double(n1[c]), 1.0 - pr, double(n0[c]) * (0.0 - pr));  // sum of (y - p) over the class
            g0 += r;
            g1 = fma(d[c], r, g1);
        }
        SmallMat<2> H;
        H(0, 0) = h00 + kL2;
        H(1, 0) = H(0, 1) = h01;
        H(1, 1) = h11 + kL2;
        if (converged) {
            double diag[2];
            lu_inverse_diag<2>(H, 2, diag);  // :138-139
            const double se = sqrt(diag[1]);
            const double zs = b1 / se;
            return {2.0 * (1.0 - normal_cdf_like_reference(fabs(zs))), b1, se, kPValue};  // :151-152
        }
        if (pass == kMaxIter) break;
        double g[2] = {g0 - kL2 * b0, g1 - kL2 * b1};  // :105
        int tr[2];
        ldlt_factor<2>(H, 2, tr);
        if (isnan(H(0, 0)) || isnan(H(1, 1))) return {nan, nan, nan, kPNa};  // :107
        ldlt_solve<2>(H, 2, tr, g);
        b0 += g[0];
        b1 += g[1];
        const double e0 = b0 - o0, e1 = b1 - o1;
        if (sqrt(fma(e1, e1, e0 * e0)) < kTol) {  // :113
            converged = true;
        } else {
            o0 = b0;
            o1 = b1;
        }
    }
    return {nan, nan, nan, kPNa};  // :120
}

}  // namespace

__global__ void __launch_bounds__(128) logit_irls_kernel(const LogitCounts* __restrict__ cnt, const BatchDev b,
                                                         const stoat_gpu_thresholds thr, LogitOut out,
                                                         uint32_t* __restrict__ general_list,
                                                         uint32_t* __restrict__ n_general) {
    const uint64_t s = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (s >= b.n_snarls) return;
    const LogitCounts c = cnt[s];
    if (c.al0 == kLogitGeneral) return;  // the general kernel writes every output of this snarl
    const double nan = __longlong_as_double(0x7ff8000000000000LL);
    const uint64_t p0 = b.snarl_path_off[s];
    const uint32_t n_used = c.used;
    const uint32_t n_full1 = n_used - c.full0 - c.half - c.third - c.tthird;
    const bool nondyadic = (c.third | c.tthird) != 0;
    // filtration_quantitative_table (src/snarl_analyzer.cpp:542-579), as quant_lean2 (quant_lean.cuh)
    const double third = 1.0 / 3.0, tthird = 2.0 / 3.0;
    const double colsum0 = double(c.full0) + 0.5 * double(c.half) + third * double(c.third) + tthird * double(c.tthird);
    const double colsum1 = double(n_full1) + 0.5 * double(c.half) + tthird * double(c.third) + third * double(c.tthird);
    bool filtered = (n_used == 0) || (c.al0 == 0) || (c.al1 == 0) || (uint64_t(n_used) < thr.min_individuals);
    bool general = false;
    if (!filtered) {
        const double total = double(n_used);  // every used row sums to 1
        if (nondyadic && uint64_t(n_used) == thr.min_haplotypes) general = true;  // borderline: exact replay
        if (total < double(thr.min_haplotypes)) filtered = true;
        const double f0 = __ddiv_rn(colsum0, total), f1 = __ddiv_rn(colsum1, total);
        const double maf0 = fmin(f0, __dsub_rn(1.0, f0)), maf1 = fmin(f1, __dsub_rn(1.0, f1));
        if (nondyadic && (fabs(maf0 - thr.maf_threshold) <= 1e-9 || fabs(maf1 - thr.maf_threshold) <= 1e-9)) general = true;
        if (!(maf0 > thr.maf_threshold && maf1 > thr.maf_threshold)) filtered = true;
    }
    if (general) {
        general_list[atomicAdd(n_general, 1u)] = uint32_t(s);
        return;
    }
    out.allele_paths[p0] = c.al0;  // quantitative_table.cpp:54-55
    out.allele_paths[p0 + 1] = c.al1;
    out.filtered[s] = filtered ? 1 : 0;
    out.n_used[s] = n_used;
    Logit2 r{nan, nan, nan, kPNa};
    if (!filtered) {
        // path 1 is the last column, which the reference drops (snarl_analyzer.cpp:676-684): d = dosage of path 0
        const double d[5] = {1.0, 0.5, third, tthird, 0.0};
        const uint32_t n1[5] = {c.c_full0, c.c_half, c.c_third, c.c_tthird,
                                c.c_used - c.c_full0 - c.c_half - c.c_third - c.c_tthird};
        const uint32_t n0[5] = {c.full0 - n1[0], c.half - n1[1], c.third - n1[2], c.tthird - n1[3], n_full1 - n1[4]};
        r = logit_irls_classes(d, n0, n1);
    }
    out.p[s] = r.p;
    out.beta[s] = r.beta;
    out.se[s] = r.se;
    out.code[s] = r.code;
}

// ---- L3: three-path snarls ---------------------------------------------------------------------------------------------
// With every sample on one or two path-haplotypes the rows of X = [1, d0, d1] (path 2 is the column the reference drops,
// snarl_analyzer.cpp:676-684) take SIX values — (1,0) (0,1) (0,0) for a sample on one path, (1/2,1/2) (1/2,0) (0,1/2) for one
// haplotype on each of two — so the regression is a function of 6 class x phenotype counts, as the two-path one is of 5.
// A warp per snarl: the class masks word by word, popcounts of (class) and (class & case), then lane 0 applies the filters
// and runs the IRLS over the six weighted points (the reference's operations per point), Wald tests of both columns, Holm.
// Anything else — a sample on three or more path-haplotypes, an empty or missing path, identical columns (the reference
// merges them) — goes to the general kernel.  (The general kernel alone took 40 of the 46 ms of a C3 step for these 5 % of
// the snarls.)
namespace {

struct Logit3 {
    double p, beta, se;
    uint8_t code;
};

__device__ Logit3 logit_irls_classes3(const uint32_t (&n0)[6], const uint32_t (&n1)[6]) {
    constexpr double kTol = 1e-6, kL2 = 1e-4, kEps = 1e-8;  // src/stats_test.hpp:109-112
    constexpr int kMaxIter = 100;
    const double nan = __longlong_as_double(0x7ff8000000000000LL);
    const double d0[6] = {1.0, 0.0, 0.0, 0.5, 0.5, 0.0}, d1[6] = {0.0, 1.0, 0.0, 0.5, 0.0, 0.5};
    double bt[3] = {0.0, 0.0, 0.0}, old[3] = {0.0, 0.0, 0.0};
    bool converged = false;
    for (int pass = 0; pass <= kMaxIter; ++pass) {
        double h[6] = {0, 0, 0, 0, 0, 0}, g[3] = {0, 0, 0};  // lower triangle (0,0) (1,0) (1,1) (2,0) (2,1) (2,2)
#pragma unroll
        for (int c = 0; c < 6; ++c) {
            const uint32_t nt = n0[c] + n1[c];
            if (!nt) continue;
            const double z = fma(d1[c], bt[2], fma(d0[c], bt[1], bt[0]));
            const double pr = 1.0 / (1.0 + exp(-z));                  // :27-29
            const double w = fmax(kEps, fmin(1.0, pr * (1.0 - pr)));   // :93
            const double sw = sqrt(w), x1 = d0[c] * sw, x2 = d1[c] * sw;  // :97-98
            const double cnt = double(nt);
            h[0] = fma(cnt, sw * sw, h[0]);
            h[1] = fma(cnt, x1 * sw, h[1]);
            h[2] = fma(cnt, x1 * x1, h[2]);
            h[3] = fma(cnt, x2 * sw, h[3]);
            h[4] = fma(cnt, x2 * x1, h[4]);
            h[5] = fma(cnt, x2 * x2, h[5]);
            const double r = fma(double(n1[c]), 1.0 - pr, double(n0[c]) * (0.0 - pr));  // sum of (y - p) over the class
            g[0] += r;
            g[1] = fma(d0[c], r, g[1]);
            g[2] = fma(d1[c], r, g[2]);
        }
        SmallMat<3> H;
        H(0, 0) = h[0] + kL2;
        H(1, 0) = H(0, 1) = h[1];
        H(1, 1) = h[2] + kL2;
        H(2, 0) = H(0, 2) = h[3];
        H(2, 1) = H(1, 2) = h[4];
        H(2, 2) = h[5] + kL2;
        if (converged) {
            double diag[3];
            lu_inverse_diag<3>(H, 3, diag);  // :138-139
            double pv[2];
            for (int i = 0; i < 2; ++i) pv[i] = 2.0 * (1.0 - normal_cdf_like_reference(fabs(bt[1 + i] / sqrt(diag[1 + i]))));  // :151-152
            // adjusted_holm (src/utils.cpp:61-87) + the smallest adjusted value, first index on ties (:169-175)
            const int lo = pv[1] < pv[0] ? 1 : 0, hi = 1 - lo;  // stable order: ties keep index 0 first
            double adj[2];
            adj[lo] = fmin(2.0 * pv[lo], 1.0);
            adj[hi] = fmax(adj[lo], fmin(pv[hi], 1.0));
            const int best = adj[1] < adj[0] ? 1 : 0;
            return {adj[best], bt[1 + best], sqrt(diag[1 + best]), kPValue};
        }
        if (pass == kMaxIter) break;
        double st[3] = {g[0] - kL2 * bt[0], g[1] - kL2 * bt[1], g[2] - kL2 * bt[2]};  // :105
        int tr[3];
        ldlt_factor<3>(H, 3, tr);
        if (isnan(H(0, 0)) || isnan(H(1, 1)) || isnan(H(2, 2))) return {nan, nan, nan, kPNa};  // :107
        ldlt_solve<3>(H, 3, tr, st);
        double nrm = 0.0;
        for (int r = 0; r < 3; ++r) {
            bt[r] += st[r];
            const double e = bt[r] - old[r];
            nrm = fma(e, e, nrm);
        }
        if (sqrt(nrm) < kTol) {  // :113
            converged = true;
        } else {
            for (int r = 0; r < 3; ++r) old[r] = bt[r];
        }
    }
    return {nan, nan, nan, kPNa};  // :120
}

}  // namespace

__global__ void __launch_bounds__(256) logit_three_path_kernel(const uint8_t* __restrict__ rows, uint32_t pitch, uint32_t n_hap,
                                                               const uint8_t* __restrict__ case_mask, const BatchDev b,
                                                               const stoat_gpu_thresholds thr, LogitOut out,
                                                               uint32_t* __restrict__ general_list,
                                                               uint32_t* __restrict__ n_general) {
    const uint32_t lane = threadIdx.x & 31;
    const uint64_t warp = (uint64_t(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const uint64_t n_warps = (uint64_t(gridDim.x) * blockDim.x) >> 5;
    const uint32_t n_words = pitch / 4, M = 0x55555555u;
    const double nan = __longlong_as_double(0x7ff8000000000000LL);
    for (uint64_t s = warp; s < b.n_snarls; s += n_warps) {
        const uint64_t p0 = b.snarl_path_off[s], p1 = b.snarl_path_off[s + 1];
        if (!logit_three_shape(b, p0, p1)) continue;
        // lane j < 3: the rows of path j (kLogitAll: no checked edge, every haplotype)
        uint32_t my_r0 = 0xFFFFFFFEu, my_r1 = 0xFFFFFFFEu;
        bool missing = false;
        if (lane < 3) {
            const uint64_t a = b.path_edge_off[p0 + lane], z = b.path_edge_off[p0 + lane + 1];
            if (z > a) {
                my_r0 = b.edge_row[a];
                my_r1 = z - a == 2 ? b.edge_row[a + 1] : my_r0;
                missing = my_r0 == kMissingL || my_r1 == kMissingL;
            }
        }
        if (__any_sync(0xffffffffu, missing)) {  // an empty path: the table has two columns, the general kernel builds it
            if (lane == 0) general_list[atomicAdd(n_general, 1u)] = uint32_t(s);
            continue;
        }
        uint32_t r0[3], r1[3];
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            r0[j] = __shfl_sync(0xffffffffu, my_r0, j);
            r1[j] = __shfl_sync(0xffffffffu, my_r1, j);
        }
        uint32_t al[3] = {0, 0, 0}, cnt[6] = {0, 0, 0, 0, 0, 0}, cas[6] = {0, 0, 0, 0, 0, 0}, bad = 0, differ = 0;
        for (uint32_t w = lane; w < n_words; w += 32) {
            const uint32_t cm = *reinterpret_cast<const uint32_t*>(case_mask + size_t(w) * 4);
            uint32_t c1[3], any[3], f[3];
#pragma unroll
            for (int j = 0; j < 3; ++j) {
                uint32_t pw;
                if (r0[j] == 0xFFFFFFFEu) {
                    pw = valid_word(w, n_hap);
                } else {
                    pw = __ldg(reinterpret_cast<const uint32_t*>(rows + size_t(r0[j]) * pitch) + w) &
                         __ldg(reinterpret_cast<const uint32_t*>(rows + size_t(r1[j]) * pitch) + w);
                }
                al[j] += __popc(pw);
                const uint32_t sft = pw >> 1;
                c1[j] = (pw ^ sft) & M;
                any[j] = (pw | sft) & M;
                f[j] = pw - (sft & M);  // 2-bit count fields
            }
            differ |= (f[0] != f[1] ? 1u : 0u) | (f[0] != f[2] ? 2u : 0u) | (f[1] != f[2] ? 4u : 0u);
            uint32_t cls[6];
            cls[0] = any[0] & ~any[1] & ~any[2];
            cls[1] = any[1] & ~any[0] & ~any[2];
            cls[2] = any[2] & ~any[0] & ~any[1];
            cls[3] = c1[0] & c1[1] & ~any[2];
            cls[4] = c1[0] & c1[2] & ~any[1];
            cls[5] = c1[1] & c1[2] & ~any[0];
            uint32_t covered = 0;
#pragma unroll
            for (int c = 0; c < 6; ++c) {
                cnt[c] += __popc(cls[c]);
                cas[c] += __popc(cls[c] & cm);
                covered |= cls[c];
            }
            bad |= (any[0] | any[1] | any[2]) & ~covered;  // a sample on three or more path-haplotypes
        }
#pragma unroll
        for (int j = 0; j < 3; ++j) al[j] = __reduce_add_sync(0xffffffffu, al[j]);
#pragma unroll
        for (int c = 0; c < 6; ++c) {
            cnt[c] = __reduce_add_sync(0xffffffffu, cnt[c]);
            cas[c] = __reduce_add_sync(0xffffffffu, cas[c]);
        }
        bad = __reduce_or_sync(0xffffffffu, bad);
        differ = __reduce_or_sync(0xffffffffu, differ);
        if (lane != 0) continue;
        if (bad != 0u || differ != 7u || al[0] == 0 || al[1] == 0 || al[2] == 0) {
            general_list[atomicAdd(n_general, 1u)] = uint32_t(s);  // thirds, identical columns, an empty column
            continue;
        }
        const uint32_t n_used = cnt[0] + cnt[1] + cnt[2] + cnt[3] + cnt[4] + cnt[5];
        // filtration_quantitative_table (src/snarl_analyzer.cpp:542-579): every sum is dyadic, hence exact
        const double colsum[3] = {double(cnt[0]) + 0.5 * double(cnt[3] + cnt[4]), double(cnt[1]) + 0.5 * double(cnt[3] + cnt[5]),
                                  double(cnt[2]) + 0.5 * double(cnt[4] + cnt[5])};
        bool filtered = n_used == 0 || uint64_t(n_used) < thr.min_individuals;
        if (!filtered) {
            const double total = double(n_used);
            int above = 0;
            for (int j = 0; j < 3; ++j) {
                const double fr = __ddiv_rn(colsum[j], total);
                if (fmin(fr, __dsub_rn(1.0, fr)) > thr.maf_threshold) ++above;
            }
            filtered = total < double(thr.min_haplotypes) || above < 2;
        }
        for (int j = 0; j < 3; ++j) out.allele_paths[p0 + j] = al[j];
        out.filtered[s] = filtered ? 1 : 0;
        out.n_used[s] = n_used;
        Logit3 r{nan, nan, nan, kPNa};
        if (!filtered) {
            uint32_t n0[6];
            for (int c = 0; c < 6; ++c) n0[c] = cnt[c] - cas[c];
            r = logit_irls_classes3(n0, cas);
        }
        out.p[s] = r.p;
        out.beta[s] = r.beta;
        out.se[s] = r.se;
        out.code[s] = r.code;
    }
}

cudaError_t launch_logit_fast(const MatrixDev& m, const uint8_t* case_mask, const BatchDev& b,
                              const stoat_gpu_thresholds& thr, LogitCounts* counts, uint32_t* general_list,
                              uint32_t* n_general, LogitOut out, int num_sms, cudaStream_t st, uint32_t* launches) {
    if (b.n_snarls == 0) return cudaSuccess;
    const uint64_t want = (b.n_snarls + 7) / 8;  // 8 warps per block
    const uint64_t cap = uint64_t(num_sms) * 8;  // 64 warps per SM
    logit_class_counts_kernel<<<unsigned(std::min(want, cap)), 256, 0, st>>>(m.rows, uint32_t(m.pitch), uint32_t(m.n_cols),
                                                                             case_mask, b, counts, general_list, n_general);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    logit_irls_kernel<<<unsigned((b.n_snarls + 127) / 128), 128, 0, st>>>(counts, b, thr, out, general_list, n_general);
    if (launches) *launches += 2;
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    logit_three_path_kernel<<<unsigned(std::min(want, cap)), 256, 0, st>>>(m.rows, uint32_t(m.pitch), uint32_t(m.n_cols), case_mask,
                                                                           b, thr, out, general_list, n_general);
    if (launches) *launches += 1;
    return cudaGetLastError();
}

}  // namespace stoat
