// kernels_quant.cu — the quantitative-trait hot path on sm_100a.
//
//   K3  dosage_moments (fast path)  process_table_quantitative + create_quantitative_table
//        (src/quantitative_table.cpp:22-104), remove_empty / filtration / combine_identical / remove_last
//        (src/snarl_analyzer.cpp:542-684) and the X^T X / X^T y build of linear_regression
//        (src/stats_test.cpp:444-458), for snarls with <= 4 paths whose samples all have row sums 1 or 2
//        (dosages in {0, 1/2, 1}) and no covariates.  Rows are TMA-staged into shared memory exactly as in
//        K1.  X^T X and every filter quantity come from popcounts of bit-sliced per-sample masks (exact
//        integers / quarters); X^T y is one dense FP64 weighted sum per regression column.
//   K4  ols_solve  one thread per snarl: LDLT / pseudo-inverse, beta, R^2, se, Student-t tail
//        (src/stats_test.cpp:398-489) from the moments K3 wrote.
//   K3G general path  any snarl (<= 64 paths, <= 16 features, covariates, thirds ...): one warp per snarl,
//        straight from global memory, lane <-> sample, regression table rebuilt in the reference's own order.
//
// For the moment path the phenotype is centred once (y' = y - mean over all samples) when it is uploaded: a
// full-rank OLS with an intercept is invariant under that shift (beta_1, se, t, R^2 unchanged) and it keeps
// RSS = y'y - 2 b'X'y + b'X'Xb well conditioned.  Snarls where that expansion would still cancel
// (R^2 > 1 - 1e-6), and rank-deficient snarls (pseudo-inverse branch: minimum-norm beta, NOT shift invariant),
// are re-done by K3G on the raw phenotype with an explicit residual pass, as the reference does (:460-463).
#include <algorithm>
#include <cstdlib>
#include <cstring>

#include "ctx.hpp"
#include "dosage_masks.cuh"
#include "ols.cuh"
#include "ptx.cuh"
#include "stage_pipeline.cuh"

namespace stoat {

constexpr uint32_t kMissingQ = 0xFFFFFFFFu;
constexpr int kBlobBytesQ = kMaxBlob16 * 16;
constexpr uint32_t kBlobHeaderQ = 8;
constexpr int kFastMaxPaths = 4;   // paths per snarl on the fast path
constexpr int kFastMaxEdges = 4;   // checked edges per path on the fast path
constexpr int kMaxPathsG = STOAT_GPU_MAX_PATHS_PER_SNARL;
constexpr int kMaxFeat = STOAT_GPU_MAX_OLS_FEATURES;

struct QuantMoments {
    double xtx[10];  // lower triangle, row-major: (0,0) (1,0) (1,1) (2,0) (2,1) (2,2) (3,0) .. (3,3)
    double xty[4];
    double syy, sy;  // over the used samples, centred phenotype
    uint32_t n_used;
    uint8_t p;      // features = 1 + regression columns (1 => the reference's undefined case)
    uint8_t state;  // 0: nothing to solve (filtered, or on the general path), 1: solve
    uint16_t pad;
};

struct K3Params {
    const uint8_t* rows;
    const SnarlRec* rec;
    const uint8_t* blob;
    const double* ys;  // centred phenotype, permuted: ys[(i % 16) * n_words + i / 16]
    uint64_t n_snarls;
    uint32_t pitch, n_hap, n_samples, stage_rows, n_stages;
    double sy_all, syy_all;
    stoat_gpu_thresholds thr;
    QuantOut out;
    QuantMoments* mom;
    uint32_t* g_list;
    uint32_t* g_count;
    const uint32_t* list;        // K3 over a list of snarls (those K3R left over); nullptr: all snarls
    const uint32_t* list_count;  // device-side length of `list`
};


template <int NP>
__device__ __forceinline__ uint32_t pick(const uint32_t (&a)[NP], int j) {
    uint32_t r = a[0];
#pragma unroll
    for (int k = 1; k < NP; ++k) r = (k == j) ? a[k] : r;
    return r;
}

__device__ __forceinline__ void push_general(const K3Params& p, uint64_t s, uint32_t lane) {
    if (lane == 0) {
        const uint32_t slot = atomicAdd(p.g_count, 1u);
        p.g_list[slot] = uint32_t(s);
    }
}

// -------------------------------------------------------------------------------------------------
// fast path for one staged snarl; returns false when the snarl must take the general path
// -------------------------------------------------------------------------------------------------
template <int NP>
__device__ bool quant_fast(const K3Params& p, const uint8_t* blob, const uint8_t* rows_s, const double* ys_s,
                           uint32_t np, uint64_t s, uint32_t lane) {
    const uint32_t p0 = *reinterpret_cast<const uint32_t*>(blob + 4);
    const uint32_t n_words = p.pitch / 4;
    // ---- path programs -> row offsets ----------------------------------------------------------
    uint32_t code[NP];
    uint32_t roff[NP][kFastMaxEdges];
    {
        uint32_t off = kBlobHeaderQ + np;
#pragma unroll
        for (int j = 0; j < NP; ++j) {
            code[j] = (uint32_t(j) < np) ? uint32_t(blob[kBlobHeaderQ + j]) : 0xFFu;
            if (code[j] != 0xFF && code[j] > uint32_t(kFastMaxEdges)) return false;
#pragma unroll
            for (int r = 0; r < kFastMaxEdges; ++r) {
                const bool live = code[j] != 0xFF && uint32_t(r) < code[j];
                roff[j][r] = live ? uint32_t(blob[off + r]) * p.pitch : 0u;
            }
            if (code[j] != 0xFF) off += code[j];
        }
    }
    auto path_word = [&](int j, uint32_t w) -> uint32_t {
        if (code[j] == 0xFF) return 0u;
        if (code[j] == 0) return valid_word(w, p.n_hap);
        uint32_t v = *reinterpret_cast<const uint32_t*>(rows_s + roff[j][0] + w * 4);
#pragma unroll
        for (int r = 1; r < kFastMaxEdges; ++r)
            if (uint32_t(r) < code[j]) v &= *reinterpret_cast<const uint32_t*>(rows_s + roff[j][r] + w * 4);
        return v;
    };
    // per-sample masks of one word (16 samples, at the even bit positions)
    struct Word {
        uint32_t c1[NP], c2[NP], a1, a2, used, bad;
    };
    auto decode = [&](uint32_t w, Word& d, uint32_t* raw) {
        uint32_t s0 = 0, s1 = 0, ovf = 0;
#pragma unroll
        for (int j = 0; j < NP; ++j) {
            const uint32_t v = path_word(j, w);
            if (raw) raw[j] = v;
            const uint32_t lo = v & 0x55555555u, hi = (v >> 1) & 0x55555555u;
            d.c1[j] = lo ^ hi;
            d.c2[j] = lo & hi;
            uint32_t c = s0 & lo;  // bit-sliced row sum += lo
            s0 ^= lo;
            uint32_t c2 = s1 & c;
            s1 ^= c;
            ovf |= c2;
            c = s0 & hi;  // += hi
            s0 ^= hi;
            c2 = s1 & c;
            s1 ^= c;
            ovf |= c2;
        }
        d.bad = ovf | (s0 & s1);  // a sample with row sum >= 3: dosages are no longer dyadic
        d.a1 = s0 & ~s1;
        d.a2 = s1 & ~s0;
        d.used = s0 | s1 | ovf;
    };

    // ---- pass 1: integer statistics ---------------------------------------------------------------
    uint32_t allele[NP], n1[NP], nh[NP], n2[NP];
    uint32_t nhh[NP * (NP - 1) / 2], diff[NP * (NP - 1) / 2];
    uint32_t n_used = 0, bad = 0;
#pragma unroll
    for (int j = 0; j < NP; ++j) allele[j] = n1[j] = nh[j] = n2[j] = 0;
#pragma unroll
    for (int q = 0; q < NP * (NP - 1) / 2; ++q) nhh[q] = diff[q] = 0;
    for (uint32_t w = lane; w < n_words; w += 32) {
        Word d;
        uint32_t raw[NP];
        decode(w, d, raw);
        bad |= d.bad;
        n_used += __popc(d.used);
        int q = 0;
#pragma unroll
        for (int j = 0; j < NP; ++j) {
            allele[j] += __popc(raw[j]);
            n1[j] += __popc(d.c1[j] & d.a1);
            nh[j] += __popc(d.c1[j] & d.a2);
            n2[j] += __popc(d.c2[j]);
#pragma unroll
            for (int l = j + 1; l < NP; ++l, ++q) {
                nhh[q] += __popc(d.c1[j] & d.c1[l] & d.a2);
                diff[q] |= (d.c1[j] ^ d.c1[l]) | (d.c2[j] ^ d.c2[l]);
            }
        }
    }
    bad = __reduce_or_sync(0xffffffffu, bad);
    if (bad) return false;
    n_used = __reduce_add_sync(0xffffffffu, n_used);
#pragma unroll
    for (int j = 0; j < NP; ++j) {
        allele[j] = __reduce_add_sync(0xffffffffu, allele[j]);
        n1[j] = __reduce_add_sync(0xffffffffu, n1[j]);
        nh[j] = __reduce_add_sync(0xffffffffu, nh[j]);
        n2[j] = __reduce_add_sync(0xffffffffu, n2[j]);
    }
#pragma unroll
    for (int q = 0; q < NP * (NP - 1) / 2; ++q) {
        nhh[q] = __reduce_add_sync(0xffffffffu, nhh[q]);
        diff[q] = __reduce_or_sync(0xffffffffu, diff[q]);
    }
    if (lane < np) p.out.allele_paths[p0 + lane] = pick<NP>(allele, int(lane));  // quantitative_table.cpp:54-55

    // ---- filters (snarl_analyzer.cpp:542-579), all quantities exact multiples of 1/2 ---------------
    int kept[NP];
    int nk = 0;
#pragma unroll
    for (int j = 0; j < NP; ++j)
        if (allele[j] > 0) {  // empty paths never become columns (quantitative_table.cpp:43-44)
#pragma unroll
            for (int k = 0; k < NP; ++k)
                if (k == nk) kept[k] = j;
            ++nk;
        }
    bool filtered = (n_used == 0) || (nk < 2) || (uint64_t(n_used) < p.thr.min_individuals);
    if (!filtered) {
        double total = 0.0;
#pragma unroll
        for (int j = 0; j < NP; ++j)
            if (allele[j] > 0) total += double(n1[j]) + 0.5 * double(nh[j]) + double(n2[j]);
        if (total < double(p.thr.min_haplotypes)) filtered = true;
        int above = 0;
#pragma unroll
        for (int j = 0; j < NP; ++j)
            if (allele[j] > 0) {
                const double colsum = double(n1[j]) + 0.5 * double(nh[j]) + double(n2[j]);
                const double freq = __ddiv_rn(colsum, total);
                const double maf = fmin(freq, __dsub_rn(1.0, freq));
                if (maf > p.thr.maf_threshold) ++above;
            }
        if (above < 2) filtered = true;
    }
    if (lane == 0) {
        p.out.filtered[s] = filtered ? 1 : 0;
        p.out.n_used[s] = n_used;
    }
    if (filtered) {
        if (lane == 0) p.mom[s].state = 0;
        return true;
    }

    // ---- combine identical columns (only with >= 3 columns), drop the last one (:620-684) ----------
    int fcol[NP], fmult[NP];  // final columns: path index, multiplicity
    int nf = 0;
    {
        bool merged[NP];
#pragma unroll
        for (int k = 0; k < NP; ++k) merged[k] = false;
#pragma unroll
        for (int a = 0; a < NP; ++a) {
            if (a >= nk || merged[a]) continue;
            int mult = 1;
            if (nk >= 3) {
#pragma unroll
                for (int b2 = a + 1; b2 < NP; ++b2) {
                    if (b2 >= nk || merged[b2]) continue;
                    // pair index of (kept[a], kept[b2]) in the upper-triangle enumeration
                    const int ja = kept[a], jb = kept[b2];
                    uint32_t dv = 1;
                    int q = 0;
#pragma unroll
                    for (int j = 0; j < NP; ++j)
#pragma unroll
                        for (int l = j + 1; l < NP; ++l, ++q)
                            if (j == ja && l == jb) dv = diff[q];
                    if (dv == 0) {
                        ++mult;
                        merged[b2] = true;
                    }
                }
            }
#pragma unroll
            for (int k = 0; k < NP; ++k)
                if (k == nf) {
                    fcol[k] = kept[a];
                    fmult[k] = mult;
                }
            ++nf;
        }
    }
    const int k_reg = nf - 1;  // remove_last_columns_quantitative_table
    QuantMoments mom;
    mom.n_used = n_used;
    mom.p = uint8_t(1 + k_reg);
    mom.state = 1;
    mom.pad = 0;
#pragma unroll
    for (int i = 0; i < 10; ++i) mom.xtx[i] = 0.0;
#pragma unroll
    for (int i = 0; i < 4; ++i) mom.xty[i] = 0.0;

    // ---- pass 2: X^T y, sums over the used samples -------------------------------------------------
    double acc[NP - 1];
#pragma unroll
    for (int a = 0; a < NP - 1; ++a) acc[a] = 0.0;
    double su = 0.0, suu = 0.0;
    const uint32_t n_unused = p.n_samples - n_used;
    const bool walk_unused = n_unused <= n_used;
    for (uint32_t w = lane; w < n_words; w += 32) {
        Word d;
        decode(w, d, nullptr);
        // 2-bit dosage fields of each regression column: 0, 1 (= 1/2), 2 (= 1)
        uint32_t field[NP - 1];
#pragma unroll
        for (int a = 0; a < NP - 1; ++a) {
            field[a] = 0;
            if (a < k_reg) {
                const uint32_t c1 = pick<NP>(d.c1, fcol[a]), c2 = pick<NP>(d.c2, fcol[a]);
                field[a] = (c1 & d.a2) | (((c1 & d.a1) | c2) << 1);
            }
        }
        const double* yw = ys_s + w;
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            const double y = yw[i * n_words];
#pragma unroll
            for (int a = 0; a < NP - 1; ++a) {
                const uint32_t bits = (field[a] >> (2 * i)) & 3u;
                // 0 -> 0.0, 1 -> 0.5 (0x3FE00000), 2 -> 1.0 (0x3FF00000)
                const int hi = bits ? int(0x3FD00000u + (bits << 20)) : 0;
                acc[a] = fma(__hiloint2double(hi, 0), y, acc[a]);
            }
        }
        // sums over used samples: walk whichever of used / unused is the smaller set
        uint32_t m = walk_unused ? (~d.used & 0x55555555u & valid_word(w, p.n_hap)) : d.used;
        while (m) {
            const int b = __ffs(m) - 1;
            m &= m - 1;
            const double y = yw[(b >> 1) * n_words];
            su += y;
            suu = fma(y, y, suu);
        }
    }
#pragma unroll
    for (int a = 0; a < NP - 1; ++a) acc[a] = warp_sum(acc[a]);
    su = warp_sum(su);
    suu = warp_sum(suu);
    if (walk_unused) {
        su = p.sy_all - su;
        suu = p.syy_all - suu;
    }
    mom.sy = su;
    mom.syy = suu;
    mom.xtx[0] = double(n_used);
    mom.xty[0] = su;
#pragma unroll
    for (int a = 0; a < NP - 1; ++a) {
        if (a >= k_reg) continue;
        const int ja = fcol[a];
        const double ma = double(fmult[a]);
        const double c_n1 = double(pick<NP>(n1, ja)), c_nh = double(pick<NP>(nh, ja)), c_n2 = double(pick<NP>(n2, ja));
        const int row = (a + 1) * (a + 2) / 2;  // start of row a+1 in the packed lower triangle
        mom.xtx[row] = ma * (c_n1 + 0.5 * c_nh + c_n2);
        mom.xtx[row + a + 1] = ma * ma * (c_n1 + 0.25 * c_nh + c_n2);
#pragma unroll
        for (int b2 = 0; b2 < NP - 1; ++b2) {
            if (b2 >= a) continue;
            const int jb = fcol[b2];
            const int lo = ja < jb ? ja : jb, hi2 = ja < jb ? jb : ja;
            uint32_t cnt = 0;
            int q = 0;
#pragma unroll
            for (int j = 0; j < NP; ++j)
#pragma unroll
                for (int l = j + 1; l < NP; ++l, ++q)
                    if (j == lo && l == hi2) cnt = nhh[q];
            mom.xtx[row + b2 + 1] = ma * double(fmult[b2]) * 0.25 * double(cnt);
        }
        mom.xty[a + 1] = ma * acc[a];
    }
    if (lane == 0) p.mom[s] = mom;
    return true;
}

#include "quant_lean.cuh"
#include "quant_reg.cuh"
#include "quant_mma.cuh"

// =================================================================================================
// K3 kernel: per-warp TMA pipelines (same structure as K1), fast path per snarl
// =================================================================================================
__global__ void __launch_bounds__(384, 1) dosage_moments_tma_kernel(const K3Params p) {
    extern __shared__ __align__(128) uint8_t smem[];
    const uint32_t n_warps = blockDim.x >> 5;
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint64_t n_items = p.list ? uint64_t(__ldg(p.list_count)) : p.n_snarls;
    if (uint64_t(blockIdx.x) * n_warps >= n_items) return;  // CTA-uniform: a short list leaves most of the grid idle
    const uint32_t stage_bytes = kBlobBytesQ + p.stage_rows * p.pitch;
    const uint32_t n_words = p.pitch / 4;
    uint64_t* full = reinterpret_cast<uint64_t*>(smem);
    const uint32_t bar_bytes = (n_warps * p.n_stages * 8 + 127) & ~127u;
    double* ys_s = reinterpret_cast<double*>(smem + bar_bytes);
    const uint32_t ys_bytes = (16u * n_words * 8u + 127) & ~127u;
    uint8_t* stages = smem + bar_bytes + ys_bytes;

    if (threadIdx.x == 0) {
        for (uint32_t i = 0; i < n_warps * p.n_stages; ++i) mbar_init(&full[i], 1);
        mbar_fence_init();
    }
    for (uint32_t i = threadIdx.x; i < 16u * n_words; i += blockDim.x) ys_s[i] = __ldg(p.ys + i);
    __syncthreads();

    auto snarl_at = [&](uint64_t k) -> uint64_t { return p.list ? uint64_t(__ldg(p.list + k)) : k; };
    const uint64_t gw = uint64_t(blockIdx.x) * n_warps + warp;
    const uint64_t stride = uint64_t(gridDim.x) * n_warps;
    uint64_t* my_full = full + warp * p.n_stages;
    uint8_t* my_stages = stages + size_t(warp) * p.n_stages * stage_bytes;
    const uint64_t policy = make_evict_first_policy();

    // prologue: fill the pipeline (lane-distributed issue, consecutive rows merged: stage_pipeline.cuh)
    {
        uint64_t k = gw;
        for (uint32_t st = 0; st < p.n_stages && k < n_items; ++st, k += stride)
            issue_stage(load_rec_word(p.rec, snarl_at(k), lane), lane, my_stages + size_t(st) * stage_bytes, &my_full[st],
                        p.rows, p.blob, p.pitch, policy);
    }

    uint32_t it = 0;
    for (uint64_t k = gw; k < n_items; k += stride, ++it) {
        const uint64_t s = snarl_at(k);
        const uint32_t stage = it % p.n_stages;
        const uint32_t parity = (it / p.n_stages) & 1u;
        const uint64_t k_next = k + uint64_t(p.n_stages) * stride;
        uint32_t next_word = (k_next < n_items) ? load_rec_word(p.rec, snarl_at(k_next), lane) : 0u;
        mbar_wait(&my_full[stage], parity);
        const uint8_t* blob = my_stages + size_t(stage) * stage_bytes;
        const uint8_t* rows_s = blob + kBlobBytesQ;
        const uint32_t np = blob[0];
        if (np) {  // np == 0: prep already put the snarl on the general list
            bool done = false;
            if (np == 2)
                done = quant_lean2(p, blob, rows_s, ys_s, s, lane);
            else if (np == 1)
                done = quant_fast<2>(p, blob, rows_s, ys_s, np, s, lane);
            else if (np == 3)
                done = quant_fast<3>(p, blob, rows_s, ys_s, np, s, lane);
            else if (np == 4)
                done = quant_fast<4>(p, blob, rows_s, ys_s, np, s, lane);
            if (!done) {
                if (lane == 0) p.mom[s].state = 0;
                push_general(p, s, lane);
            }
        } else if (lane == 0) {
            p.mom[s].state = 0;
        }
        __syncwarp();
        asm volatile("" : "+r"(next_word));
        if (k_next < n_items)  // WAR on a read-only stage: no proxy fence (see issue_stage)
            issue_stage(next_word, lane, my_stages + size_t(stage) * stage_bytes, &my_full[stage], p.rows, p.blob,
                        p.pitch, policy);
    }
}

// =================================================================================================
// K4: solve
// =================================================================================================
// state 1: moments written by K3 / K3M3; state 2: two-path snarl whose class counts and sums K3M left in `part`
// (n_slices slices, added here) — filters and moments are then built in registers (lean_finish).
//
// A block takes kSolveChunk consecutive snarls.  State-2 snarls (the bulk) are solved thread-per-snarl by solve_two, the
// 2 x 2 case of xtx_inverse written out on scalars (no local-memory matrices, no loops).  State-1 snarls are only
// collected into a shared-memory list and solved afterwards, densely packed, by the general small-matrix code: mixed
// in the same warps they made ~80 % of the warps run both code paths (ncu r1f_k4.txt: 20 of 32 lanes active, 33 % of
// the stalls waiting for instructions).
constexpr int kSolveThreads = 256;
constexpr int kSolveChunk = 1024;

struct SolveArgs {
    QuantOut out;
    uint32_t* g_list;
    uint32_t* g_count;
    uint32_t defer;
    double* tail_t;   // [S] prefactor / a of the fraction's side
    double* tail_x;   // [2][S] argument x, then the state d = h the fraction starts from (IbetaCf::init)
    uint32_t* tail_df;
    uint64_t n_snarls;
};

__device__ __forceinline__ void solve_hand_on(const SolveArgs& a, uint64_t s) {
    if (a.defer & 1u) {
        a.out.code[s] = kPDeferred;
        a.out.filtered[s] = 0;
    } else {
        const uint32_t slot = atomicAdd(a.g_count, 1u);
        a.g_list[slot] = uint32_t(s);
    }
}

// the tail of linear_regression (src/stats_test.cpp:460-489) as in ols_statistics (ols.cuh), except that the
// Student-t tail — a continued fraction of 5 .. hundreds of iterations — is left to K4b (student_tail_kernel)
__device__ __forceinline__ void solve_store(const SolveArgs& a, uint64_t s, int p, uint32_t n_used, double beta1,
                                            double inv11, double rss, double tss) {
    int df_res = int(n_used) - p + 1;  // :467 (sic)
    if (df_res < 1) df_res = 1;        // :468
    const double mse = rss / double(df_res);
    const double se = sqrt(inv11 * mse);
    const double t = beta1 / se;
    a.out.r2[s] = 1.0 - (rss / tss);
    a.out.beta[s] = beta1;
    a.out.se[s] = se;
    a.out.code[s] = kPValue;
    if (isnan(t) || isinf(t)) {
        a.out.p[s] = 1.0;  // :480-482
        return;
    }
    // set-up of student_t_upper_tail (device_math.cuh) here, where every lane has one: the prefactor
    // x^a (1-x)^b / B(a, b) and the argument / side of the continued fraction, which K4b then only iterates
    const double df = double(df_res), t2 = t * t;
    if (t2 == 0.0) {
        a.out.p[s] = 1.0;  // 2 * 0.5
        return;
    }
    const double aa = 0.5 * df, bb = 0.5;
    const double z = t2 / (df + t2);  // 1 - x
    const double log_front = -aa * log1p(t2 / df) + bb * log(z) - log_beta_half(aa);
    const bool upper_side = z > 1.5 / (aa + 2.5);
    // ... and the first step of ibeta_cf (IbetaCf::init) plus the division by a of the closing formula, so that a
    // K4b refill and a K4b finish — executed with a handful of lanes active — contain no division at all
    const double x = upper_side ? df / (df + t2) : z;
    const double ca = upper_side ? aa : bb, cb = upper_side ? bb : aa;
    double d0 = 1.0 - (ca + cb) * x / (ca + 1.0);
    if (fabs(d0) < 1e-300) d0 = 1e-300;
    d0 = 1.0 / d0;
    a.tail_t[s] = exp(log_front) / ca;
    a.tail_x[s] = x;
    a.tail_x[a.n_snarls + s] = d0;
    a.tail_df[s] = uint32_t(df_res) | (upper_side ? 0x80000000u : 0u);
}

// general case: X^T X up to 4 x 4 (ols.cuh)
__device__ __noinline__ void solve_general(const SolveArgs& a, uint64_t s, const QuantMoments& m) {
    const int p = m.p;
    const double nan = __longlong_as_double(0x7ff8000000000000LL);
    if (p < 2) {  // no regression column left: the reference indexes past the end (stats_test.cpp:487-489)
        a.out.p[s] = a.out.r2[s] = a.out.beta[s] = a.out.se[s] = nan;
        a.out.code[s] = kPUndef;
        return;
    }
    SmallMat<4> X, inv;
    int q = 0;
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j <= i; ++j, ++q) {
            X(i, j) = m.xtx[q];
            X(j, i) = m.xtx[q];
        }
    if (xtx_inverse<4>(X, p, inv)) {
        // rank deficient: the pseudo-inverse picks the minimum-norm solution, which is NOT invariant under the
        // centring of y used by the moment path -> redo on the general path with the raw phenotype
        solve_hand_on(a, s);
        return;
    }
    double beta[4];
    for (int i = 0; i < p; ++i) {
        double b = 0.0;
        for (int j = 0; j < p; ++j) b += inv(i, j) * m.xty[j];
        beta[i] = b;
    }
    // rss = y'y - 2 b'X'y + b'(X'X)b ; tss = y'y - (sum y)^2 / n
    double bxy = 0.0, bxxb = 0.0;
    for (int i = 0; i < p; ++i) {
        bxy += beta[i] * m.xty[i];
        double r = 0.0;
        for (int j = 0; j < p; ++j) r += X(i, j) * beta[j];
        bxxb += beta[i] * r;
    }
    const double rss = m.syy - 2.0 * bxy + bxxb;
    const double tss = m.syy - m.sy * m.sy / double(m.n_used);
    if (!(rss > 1e-6 * tss)) {
        // expansion not trustworthy (near-perfect fit): redo with explicit residuals on the general path
        solve_hand_on(a, s);
        return;
    }
    solve_store(a, s, p, m.n_used, beta[1], inv(1, 1), rss, tss);
}

// X^T X = [[n, c], [c, q]]: xtx_inverse<N>(.., 2, ..) of ols.cuh step by step on scalars — LDL^T with the pivot on the
// larger diagonal entry (n >= q for dosages in [0, 1]; otherwise the general code), the two unit-vector solves, beta,
// rss by the same expansion.  Returns false when the general code must take the snarl.
__device__ __forceinline__ bool solve_two(const SolveArgs& a, uint64_t s, double n, double c, double q, double xty0,
                                          double xty1, double syy, double sy, uint32_t n_used) {
    if (fabs(q) > fabs(n)) return false;
    const double d0 = n;
    if (!(fabs(d0) >= 1e-10)) return false;
    const double l = c / d0;
    const double d1 = q - l * (d0 * l);
    if (!(fabs(d1) >= 1e-10)) return false;  // rank deficient: decided by the general code
    const double i11 = 1.0 / d1;
    const double i10 = (0.0 - l) / d1;
    const double i00 = 1.0 / d0 - l * i10;
    const double i01 = 0.0 - l * i11;
    const double beta0 = i00 * xty0 + i01 * xty1;
    const double beta1 = i10 * xty0 + i11 * xty1;
    const double bxy = beta0 * xty0 + beta1 * xty1;
    const double r0 = n * beta0 + c * beta1, r1 = c * beta0 + q * beta1;
    const double bxxb = beta0 * r0 + beta1 * r1;
    const double rss = syy - 2.0 * bxy + bxxb;
    const double tss = syy - sy * sy / double(n_used);
    if (!(rss > 1e-6 * tss)) {
        solve_hand_on(a, s);
        return true;
    }
    solve_store(a, s, 2, n_used, beta1, i11, rss, tss);
    return true;
}

__global__ void __launch_bounds__(kSolveThreads) ols_solve_kernel(const QuantMoments* __restrict__ mom, uint64_t n_snarls,
                                                                  const K3Params kp, const LeanPartials part,
                                                                  uint32_t n_slices,
                                                                  const uint64_t* __restrict__ snarl_path_off,
                                                                  QuantOut out, uint32_t* g_list, uint32_t* g_count,
                                                                  uint32_t defer, double* __restrict__ tail_t,
                                                                  double* __restrict__ tail_x,
                                                                  uint32_t* __restrict__ tail_df) {
    // defer (sample-sharded runs): a snarl that needs more than the summed two-path partials — any other shape, a
    // rank-deficient or near-perfect fit — is not sent to the general path (this GPU holds only some of the columns)
    // but marked STOAT_P_DEFERRED; the host replays those snarls with all columns (stoat_b200/shard.py).
    __shared__ uint32_t later[kSolveChunk];
    __shared__ uint32_t n_later;
    const SolveArgs a{out, g_list, g_count, defer, tail_t, tail_x, tail_df, n_snarls};
    if (threadIdx.x == 0) n_later = 0u;
    __syncthreads();
    const uint64_t base = uint64_t(blockIdx.x) * kSolveChunk;
    for (uint32_t it = 0; it < uint32_t(kSolveChunk / kSolveThreads); ++it) {
        const uint64_t s = base + uint64_t(it) * kSolveThreads + threadIdx.x;
        const uint8_t state = s < n_snarls ? mom[s].state : uint8_t(0xFF);
        bool general = state == 1;
        if (s < n_snarls) {
            // the outputs of a snarl nobody solves (filtered, or handed on) — this kernel visits every snarl, so it
            // replaces six memsets per step; whoever solves the snarl (below, K4b, the general path) writes later
            const double na = __longlong_as_double(-1LL);
            out.p[s] = na;
            out.r2[s] = na;
            out.beta[s] = na;
            out.se[s] = na;
            out.code[s] = kPNa;
            tail_t[s] = na;
        }
        if (state == 2) {
            LeanSums t{};
            for (uint32_t sl = 0; sl < n_slices; ++sl) {
                const uint32_t* pi = part.i + size_t(sl) * 7 * part.stride + s;
                const double* pd = part.d + size_t(sl) * 3 * part.stride + s;
                t.al0 += pi[0];
                t.al1 += pi[part.stride];
                t.n_used += pi[2 * part.stride];
                t.n_full0 += pi[3 * part.stride];
                t.n_half += pi[4 * part.stride];
                t.n_third += pi[5 * part.stride];
                t.n_tthird += pi[6 * part.stride];
                t.xty1 += pd[0];
                t.su_un += pd[part.stride];
                t.suu_un += pd[2 * part.stride];
            }
            QuantMoments m;
            const int r = lean_finish(kp, s, uint32_t(snarl_path_off[s]), t, m);
            if (r == 2) solve_hand_on(a, s);
            // false: rank deficient (or q > n, impossible for dosage columns) — what xtx_inverse decides there
            if (r == 1 && !solve_two(a, s, m.xtx[0], m.xtx[1], m.xtx[2], m.xty[0], m.xty[1], m.syy, m.sy, m.n_used)) {
                const QuantMoments copy = m;  // only the copy's address escapes: m itself stays in registers
                solve_general(a, s, copy);
            }
        } else if (state != 1 && state != 0xFF) {
            if (defer & 1u) solve_hand_on(a, s);
        }
        const uint32_t votes = __ballot_sync(0xffffffffu, general);
        if (votes != 0u) {
            const uint32_t lane = threadIdx.x & 31u;
            uint32_t at = 0;
            if (lane == 0) at = atomicAdd(&n_later, uint32_t(__popc(votes)));
            at = __shfl_sync(0xffffffffu, at, 0);
            if (general) later[at + __popc(votes & ((1u << lane) - 1u))] = uint32_t(s - base);
        }
    }
    __syncthreads();
    const uint32_t n_gen = n_later;
    for (uint32_t i = threadIdx.x; i < n_gen; i += kSolveThreads) {
        const uint64_t s = base + later[i];
        const QuantMoments m = mom[s];
        solve_general(a, s, m);
    }
}

// K4b: p = 2 * P(T_df > |t|) (stats_test.cpp:484) for every snarl K4 left a prefactor for (tail_t[s] is NaN otherwise).
// One lane = one running continued fraction (student_t_upper_tail of device_math.cuh, operation for operation); a lane
// whose fraction has converged takes the next snarl, so the lanes of a warp stay in the common loop body
// (thread-per-snarl: 6.5 of 32 lanes active, ncu r1e_k4.txt).  The fractions are short (mean 7 iterations, median 5 on
// the C3 data), so some lane refills in nearly every iteration: the refill must not touch global memory.  Each warp
// therefore takes chunks of kTailChunk snarls from the shared counter, loads them coalesced, drops the snarls with
// nothing to evaluate and keeps the rest as a queue in shared memory; a refill is a ballot, a rank and three LDS.
// The expensive set-up (log1p, log, exp, ln B) was done by K4 with all lanes busy.
constexpr int kTailChunk = 64;
constexpr int kTailWarps = 4;
__global__ void __launch_bounds__(kTailWarps * 32) student_tail_kernel(const double* __restrict__ tail_t,
                                                                       const double* __restrict__ tail_x,
                                                                       const uint32_t* __restrict__ tail_df,
                                                                       uint32_t n_snarls, uint32_t* __restrict__ next,
                                                                       double* __restrict__ p_out) {
    __shared__ double q_front[kTailWarps][kTailChunk], q_x[kTailWarps][kTailChunk], q_d[kTailWarps][kTailChunk];
    __shared__ uint32_t q_df[kTailWarps][kTailChunk], q_idx[kTailWarps][kTailChunk];
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t lt = (1u << lane) - 1u;
    bool active = false, drained = false;  // drained: warp-uniform
    uint32_t avail = 0, taken = 0;         // the queue holds items [taken, avail) (warp-uniform)
    uint32_t snarl = 0;
    IbetaCf cf;
    double front = 0.0;
    bool upper_side = false;
    for (;;) {
        const uint32_t need = __ballot_sync(0xffffffffu, !active);
        if (need) {
            while (taken == avail && !drained) {  // queue empty: next chunk
                uint32_t base = 0;
                if (lane == 0) base = atomicAdd(next, uint32_t(kTailChunk));
                base = __shfl_sync(0xffffffffu, base, 0);
                __syncwarp();
                avail = taken = 0;
                if (base >= n_snarls) {
                    drained = true;
                    break;
                }
#pragma unroll
                for (int k = 0; k < kTailChunk / 32; ++k) {
                    const uint32_t w = base + 32u * k + lane;
                    double f = __longlong_as_double(0x7ff8000000000000LL);
                    if (w < n_snarls) f = tail_t[w];
                    const bool has = !isnan(f);
                    const uint32_t votes = __ballot_sync(0xffffffffu, has);
                    if (has) {
                        const uint32_t at = avail + __popc(votes & lt);
                        q_front[warp][at] = f;
                        q_x[warp][at] = tail_x[w];
                        q_d[warp][at] = tail_x[size_t(n_snarls) + w];
                        q_df[warp][at] = tail_df[w];
                        q_idx[warp][at] = w;
                    }
                    avail += __popc(votes);
                }
                __syncwarp();
            }
            const uint32_t left = avail - taken;
            if (!active) {
                const uint32_t r = __popc(need & lt);
                if (r < left) {
                    const uint32_t at = taken + r;
                    front = q_front[warp][at];
                    const uint32_t d = q_df[warp][at];
                    const double a = 0.5 * double(d & 0x7FFFFFFFu);
                    upper_side = (d >> 31) != 0u;
                    cf.init_from(upper_side ? a : 0.5, upper_side ? 0.5 : a, q_x[warp][at], q_d[warp][at]);
                    snarl = q_idx[warp][at];
                    active = true;
                }
            }
            taken += min(uint32_t(__popc(need)), left);
        }
        if (!__any_sync(0xffffffffu, active)) {
            if (drained && taken == avail) break;
            continue;
        }
        if (active && cf.step()) {
            // 2 * (0.5 * front * h / a) or 2 * 0.5 * (1 - front * h / a), front / a taken by K4
            p_out[snarl] = upper_side ? front * cf.h : 1.0 - front * cf.h;
            active = false;
        }
    }
}

#include "quant_general.cuh"

// Rows too wide for K3's shared-memory phenotype table (16 x pitch / 4 doubles: 200 KB at 50,000 haplotypes): what K3M
// and its siblings left on the slow list goes to the general kernel instead of K3.
__global__ void slow_list_to_general_kernel(const uint32_t* __restrict__ slow_list, const uint32_t* __restrict__ slow_count,
                                            QuantMoments* __restrict__ mom, uint32_t* __restrict__ g_list,
                                            uint32_t* __restrict__ g_count) {
    const uint32_t n = *slow_count;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t s = slow_list[i];
        mom[s].state = 0;
        g_list[atomicAdd(g_count, 1u)] = s;
    }
}


// =================================================================================================
// host side
// =================================================================================================
struct K3Config {
    uint32_t n_warps, n_stages, stage_rows, smem_bytes;
};
static bool choose_k3_config(uint32_t pitch, uint32_t stage_rows, K3Config& cfg) {
    const uint32_t budget = 220 * 1024;
    if (stage_rows == 0) stage_rows = 1;
    const uint32_t stage_bytes = kBlobBytesQ + stage_rows * pitch;
    const uint32_t ys_bytes = (16u * (pitch / 4) * 8u + 127) & ~127u;
    if (ys_bytes + 2 * stage_bytes + 1024 > budget) return false;
    const uint32_t total_stages = (budget - ys_bytes - 1024) / stage_bytes;
    // as in K1: shared memory goes to warps first (2 stages each), the compute chain of a warp limits the rate
    uint32_t n_stages = 2;
    if (const char* e = std::getenv("STOAT_GPU_K3_STAGES")) {
        const int v = std::atoi(e);
        if (v >= 2 && v <= 4 && uint32_t(v) <= total_stages) n_stages = uint32_t(v);
    }
    uint32_t n_warps = total_stages / n_stages;
    if (n_warps > 12) n_warps = 12;
    if (n_warps == 0) return false;
    cfg.n_warps = n_warps;
    cfg.n_stages = n_stages;
    cfg.stage_rows = stage_rows;
    const uint32_t bar_bytes = (n_warps * n_stages * 8 + 127) & ~127u;
    cfg.smem_bytes = bar_bytes + ys_bytes + n_warps * n_stages * stage_bytes;
    return true;
}

// ---- K3R launch plan ---------------------------------------------------------------------------------
struct K3RPlan {
    int wpt = 0;  // 0: K3R does not apply
    uint32_t threads = 0, n_slices = 0, slice_words = 0, n_ring = 0, smem_bytes = 0, ctas_per_sm = 0;
};
using K3RKernel = void (*)(const K3RParams);
static K3RKernel k3r_kernel(int wpt) {
    return wpt == 1 ? dosage_moments_reg_kernel<1> : wpt == 2 ? dosage_moments_reg_kernel<2> : dosage_moments_reg_kernel<4>;
}
static int env_int(const char* name, int dflt) {
    const char* e = std::getenv(name);
    return e ? std::atoi(e) : dflt;
}
static void k3r_shape(uint32_t pitch, K3RPlan& pl) {
    const uint32_t n_words = pitch / 4;
    int wpt = env_int("STOAT_GPU_K3R_WPT", 2);
    if (wpt != 1 && wpt != 2 && wpt != 4) wpt = 2;
    const uint32_t per_cta = uint32_t(kRegMaxWarps) * 32u * uint32_t(wpt);
    pl.wpt = wpt;
    pl.n_slices = (n_words + per_cta - 1) / per_cta;
    pl.slice_words = (((n_words + pl.n_slices - 1) / pl.n_slices) + 3u) & ~3u;
    pl.threads = (((pl.slice_words + wpt - 1) / wpt) + 31u) & ~31u;
}
uint32_t quant_reg_slices(uint64_t pitch) {
    if (env_int("STOAT_GPU_K3R", 0) == 0 || pitch == 0) return 0;
    K3RPlan pl;
    k3r_shape(uint32_t(pitch), pl);
    return pl.n_slices;
}
static bool k3r_plan(uint32_t pitch, uint32_t stage_rows, K3RPlan& pl) {
    if (env_int("STOAT_GPU_K3R", 0) == 0) return false;
    if (stage_rows == 0) stage_rows = 1;
    k3r_shape(pitch, pl);
    const uint32_t stage_bytes = kStageBlobBytes + stage_rows * pl.slice_words * 4;
    K3RKernel fn = k3r_kernel(pl.wpt);
    const uint32_t budget = 224 * 1024;
    if (cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, int(budget)) != cudaSuccess) return false;
    int occ = 0;  // CTAs per SM that registers / threads allow with the smallest ring
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fn, int(pl.threads), 384 + 2 * stage_bytes) != cudaSuccess ||
        occ < 1)
        return false;
    const int want = env_int("STOAT_GPU_K3R_CTAS", 0);
    if (want >= 1 && want < occ) occ = want;
    const uint32_t per_cta = (budget - 1024) / uint32_t(occ) - 1024;  // 1 KB per CTA is reserved by the driver
    uint32_t ring = (per_cta - 384) / stage_bytes;
    const int ring_env = env_int("STOAT_GPU_K3R_RING", 0);
    const uint32_t ring_max = ring_env >= 2 ? uint32_t(ring_env) : 8u;
    if (ring > ring_max) ring = ring_max;
    if (ring < uint32_t(kRegLag) + 1) return false;
    pl.n_ring = ring;
    pl.smem_bytes = (((2 * ring + 2 * kRegBufs) * 8 + 127) & ~127u) + ring * stage_bytes;
    pl.ctas_per_sm = uint32_t(occ);
    return true;
}

// ---- K3M launch plan ---------------------------------------------------------------------------------
struct K3MPlan {
    uint32_t n_slices = 0, rounds_per_slice = 0, n_rounds = 0, smem_bytes = 0;
};
static bool k3m_plan(uint32_t pitch, uint64_t n_hap, K3MPlan& pl) {
    if (env_int("STOAT_GPU_K3M", 1) == 0 || pitch == 0 || n_hap >= 65536) return false;
    const uint32_t n_rounds = (pitch / 16 + 3) / 4;
    pl.n_rounds = n_rounds;
    pl.n_slices = (n_rounds + kMmaMaxRounds - 1) / kMmaMaxRounds;
    pl.rounds_per_slice = (n_rounds + pl.n_slices - 1) / pl.n_slices;
    pl.smem_bytes = pl.rounds_per_slice * 2048;
    return true;
}
// true: launch_quantitative_impl takes the moments path, whose K4 initialises the outputs itself
bool quant_fast_path(uint64_t pitch, uint64_t n_hap, uint32_t max_stage_rows, int variant, uint32_t n_covar) {
    K3Config cfg{};
    K3MPlan pm;
    return variant == 0 && n_covar == 0 &&
           (choose_k3_config(uint32_t(pitch), max_stage_rows, cfg) || k3m_plan(uint32_t(pitch), n_hap, pm));
}
size_t quant_yslices_bytes(uint64_t pitch) { return size_t((pitch / 16 + 3) / 4) * 2048; }
// slices of the haplotype range the two-path kernel uses (sizes the partial-sum buffers); 0: neither K3M nor K3R
uint32_t quant_lean_slices(uint64_t pitch, uint64_t n_hap, uint32_t stage_rows) {
    K3MPlan pm;
    (void)stage_rows;
    if (k3m_plan(uint32_t(pitch), n_hap, pm)) return pm.n_slices;
    return quant_reg_slices(pitch);
}

// K4b over every snarl of the batch (tail_t[s] NaN: nothing to do); counters[7] is the shared work counter
static cudaError_t launch_student_tail(const double* tail_t, const double* tail_x, const uint32_t* tail_df,
                                       const BatchDev& b, double* p_out, int num_sms, cudaStream_t st) {
    cudaError_t e = cudaMemsetAsync(b.counters + 7, 0, 4, st);
    if (e != cudaSuccess) return e;
    // every warp ends with a drain of ~25 iterations in which its lanes run dry one by one, so the grid is as small as
    // keeps the FP64 pipe busy, not as large as fits (measured, blocks per SM 1/2/3/4/6/8: 472/316/271/251/243/261 us)
    static const int per_sm = env_int("STOAT_GPU_K4B_BLOCKS", 6);
    const uint64_t want = (b.n_snarls + kTailWarps * kTailChunk - 1) / (kTailWarps * kTailChunk);
    const uint64_t cap = uint64_t(num_sms) * uint64_t(per_sm < 1 ? 1 : per_sm);
    const unsigned blocks = unsigned(want < cap ? want : cap);
    student_tail_kernel<<<blocks, kTailWarps * 32, 0, st>>>(tail_t, tail_x, tail_df, uint32_t(b.n_snarls), b.counters + 7, p_out);
    return cudaGetLastError();
}

// scratch owned by the caller: mom [n_snarls], ys permuted, sums; K3R: part_i [slices][7][S], part_d [slices][3][S],
// slow_list [S]; counters[3] = length of the slow list
#include "quant_covar.cuh"

cudaError_t launch_quantitative_impl(const MatrixDev& m, const double* y_raw, const double* ys_perm,
                                     double sy_all, double syy_all, const double* covar, uint32_t n_covar,
                                     const BatchDev& b, uint32_t max_stage_rows, const stoat_gpu_thresholds& thr,
                                     QuantOut out, QuantMoments* mom, uint32_t* part_i, double* part_d,
                                     uint32_t* slow_list, uint4* lean_rec, const uint8_t* yslices, const double* yc,
                                     double y_scale, int num_sms, int variant, cudaStream_t st, uint32_t* launches,
                                     cudaEvent_t ev_table_done, int phase, double* tail_t, double* tail_x,
                                     uint32_t* tail_df, uint64_t stage_gen, uint64_t* lean_key, void* cov_mom,
                                     const double* cov_prefix, uint32_t cov_rows) {
    // phase 0: the whole step.  Sample-sharded runs split it around the all-reduce of the partial sums:
    // phase 1 = lean records + K3M only, phase 2 = K4 on the summed partials with everything else deferred to the host.
    if (b.n_snarls == 0) return cudaSuccess;
    const uint32_t pitch = uint32_t(m.pitch);
    const uint32_t n_samples = uint32_t(m.n_cols / 2);
    K3Config cfg{};
    // counters[0] already holds the number of snarls prep sent to the general list.  K3 needs the permuted phenotype in
    // shared memory (32 x pitch bytes); where that does not fit (> ~37,000 haplotypes) the moments path still runs on
    // K3M, whose slices bound its shared memory, and the snarls K3M leaves over take the general kernel.
    const bool k3_ok = choose_k3_config(pitch, max_stage_rows, cfg);
    K3MPlan pm_probe;
    const bool k3m_ok = part_i && part_d && slow_list && lean_rec && yslices && yc && k3m_plan(pitch, m.n_cols, pm_probe);
    const bool fast = (variant == 0) && (n_covar == 0) && (k3_ok || k3m_ok);
    cudaError_t e;
    GParams g{};
    g.rows = m.rows;
    g.pitch = pitch;
    g.n_hap = uint32_t(m.n_cols);
    g.n_samples = n_samples;
    g.y = y_raw;
    g.covar = covar;
    g.n_covar = n_covar;
    g.b = b;
    g.thr = thr;
    g.out = out;
    if (fast) {
        K3Params p{};
        p.rows = m.rows;
        p.rec = b.rec;
        p.blob = b.blob;
        p.ys = ys_perm;
        p.n_snarls = b.n_snarls;
        p.pitch = pitch;
        p.n_hap = uint32_t(m.n_cols);
        p.n_samples = n_samples;
        p.stage_rows = cfg.stage_rows;
        p.n_stages = cfg.n_stages;
        p.sy_all = sy_all;
        p.syy_all = syy_all;
        p.thr = thr;
        p.out = out;
        p.mom = mom;
        p.g_list = b.big_list;
        p.g_count = b.counters;
        if (k3_ok) {
            static thread_local int k3_dev = -1;
            static thread_local uint32_t k3_smem = 0;
            int dev_now = 0;
            cudaGetDevice(&dev_now);
            if (dev_now != k3_dev || cfg.smem_bytes != k3_smem) {
                e = cudaFuncSetAttribute(dosage_moments_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         int(cfg.smem_bytes));
                if (e != cudaSuccess) return e;
                k3_dev = dev_now;
                k3_smem = cfg.smem_bytes;
            }
        }
        K3RPlan pl;
        LeanPartials part{part_i, part_d, b.n_snarls};
        uint32_t n_slices = 0;
        K3MPlan pm;
        if (phase == 2) {
            if (!k3m_plan(pitch, m.n_cols, pm)) return cudaErrorNotSupported;
            ols_solve_kernel<<<unsigned((b.n_snarls + kSolveChunk - 1) / kSolveChunk), kSolveThreads, 0, st>>>(
                mom, b.n_snarls, p, part, pm.n_slices, b.snarl_path_off, out, b.big_list, b.counters, 1u, tail_t, tail_x, tail_df);
            if ((e = cudaGetLastError()) != cudaSuccess) return e;
            if ((e = launch_student_tail(tail_t, tail_x, tail_df, b, out.p, num_sms, st)) != cudaSuccess) return e;
            if (launches) *launches += 2;
            return cudaSuccess;
        }
        if (phase == 1 && !(part_i && part_d && slow_list && lean_rec && yslices && yc && k3m_plan(pitch, m.n_cols, pm)))
            return cudaErrorNotSupported;
        if (part_i && part_d && slow_list && lean_rec && yslices && yc && k3m_plan(pitch, m.n_cols, pm)) {
            // K3M over every snarl (two-path snarls -> exact integer partial sums), then K3 over the slow list
            K3MParams q{};
            q.k = p;
            q.lean_rec = lean_rec;
            q.yslices = yslices;
            q.yc = yc;
            q.y_scale = y_scale;
            q.rounds_per_slice = pm.rounds_per_slice;
            q.n_rounds = pm.n_rounds;
            q.prefetch = env_int("STOAT_GPU_K3M_PREFETCH", 0) ? 1u : 0u;  // measured slower on B200 (1.54 vs 1.24 ms): off
            q.slow_list = slow_list;
            q.slow_count = b.counters + 3;
            q.part = part;
            n_slices = pm.n_slices;
            if ((e = cudaMemsetAsync(b.counters + 3, 0, 8, st)) != cudaSuccess) return e;
            // three-path snarls: K3M3 when one CTA row covers the haplotype range and the step is not sample-sharded
            // (K3M3 writes moments, not partial sums); their records live behind the two-path records
            // beyond one slice (> 12,288 haplotypes) their slice table stays in global memory
            const bool wide_table = pm.n_slices > 1;
            const uint32_t smem_small = wide_table ? 0u : pm.smem_bytes;
            q.table_in_global = wide_table ? 1u : 0u;
            const bool three = phase == 0 && env_int("STOAT_GPU_K3M3", 1) != 0;
            uint4* rec3 = three ? lean_rec + b.n_snarls : nullptr;
            // ... and two-path snarls with dosages of 1/3, 2/3 (K3M2x), under the same conditions: [S][2] more records
            const bool thirds = phase == 0 && env_int("STOAT_GPU_K3M2X", 1) != 0;
            q.rec2x = thirds ? lean_rec + 3 * b.n_snarls : nullptr;
            q.count2x = b.counters + 6;
            q.count2x_dyn = b.counters + 4;
            // the records depend on the staged paths and on which kernels take part, not on the phenotype or the
            // matrix: a caller that runs many steps on one staging (lean_key) builds them once.  What K3M appends to
            // the thirds list during a step is counted apart (counters[4], zeroed with the slow list's length)
            const uint64_t key = (stage_gen << 3) | 4u | (three ? 2u : 0u) | (thirds ? 1u : 0u);
            const bool cached = lean_key != nullptr && *lean_key == key;
            if (!cached) {
                if ((three || thirds) && (e = cudaMemsetAsync(b.counters + 5, 0, 8, st)) != cudaSuccess) return e;
                lean_rec_kernel<<<unsigned((b.n_snarls + 255) / 256), 256, 0, st>>>(b.rec, b.blob, b.n_snarls, lean_rec,
                                                                                    rec3, b.counters + 5, q.rec2x, q.count2x);
                if ((e = cudaGetLastError()) != cudaSuccess) return e;
                if (lean_key) *lean_key = key;
                if (launches) *launches += 1;
            }
            // K3M variant: the register kernel, or the cp.async ring kernel (STOAT_GPU_K3M_RING=<warps>x<stages>, one CTA per
            // SM).  Measured at C3 on one box: register kernel 1.26 ms; ring 8x2 1.42, 10x2 1.34, 11x2 1.23, 6x3 1.90, 7x3
            // 1.65 ms - the time follows the number of warps, not the way the rows arrive: K3M is bound by the issue rate of
            // its ~138 instructions per word step (the bare access patterns reach 5.4-6.8 TB/s, scripts/micro/k3m_pattern.cu)
            const int ring_cfg = [] {
                const char* v = std::getenv("STOAT_GPU_K3M_RING");
                if (!v || !std::strcmp(v, "0")) return 0;
                if (!std::strcmp(v, "8x2")) return 82;
                if (!std::strcmp(v, "10x2")) return 102;
                if (!std::strcmp(v, "11x2")) return 112;
                if (!std::strcmp(v, "6x3")) return 63;
                if (!std::strcmp(v, "7x3")) return 73;
                return 0;
            }();
            const uint64_t n_tiles = (b.n_snarls + kMmaTile - 1) / kMmaTile;
            int dev_now = 0;
            cudaGetDevice(&dev_now);
            K3MParams q_main = q;
            // With several haplotype slices a snarl's thirds may show in one slice only, so K3M does not append to the
            // thirds list itself (the slices would race on the snarl's state): K4 sends such a snarl to the general list
            if (wide_table) q_main.rec2x = nullptr;
            auto launch_ring = [&](auto kernel, int warps, int stages) -> cudaError_t {
                const uint32_t smem = pm.smem_bytes + uint32_t(warps * stages) * kRingStageBytes;
                static thread_local int ring_dev = -1;
                static thread_local uint32_t ring_smem = 0;
                if (dev_now != ring_dev || smem != ring_smem) {
                    cudaError_t e2 = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
                    if (e2 != cudaSuccess) return e2;
                    ring_dev = dev_now;
                    ring_smem = smem;
                }
                uint64_t ctas = uint64_t(num_sms) / pm.n_slices;
                if (ctas < 1) ctas = 1;
                if (ctas > (n_tiles + warps - 1) / warps) ctas = (n_tiles + warps - 1) / warps;
                kernel<<<dim3(unsigned(ctas), pm.n_slices), warps * 32, smem, st>>>(q_main);
                return cudaGetLastError();
            };
            if (ring_cfg == 82) e = launch_ring(dosage_moments_mma_ring_kernel<8, 2>, 8, 2);
            else if (ring_cfg == 102) e = launch_ring(dosage_moments_mma_ring_kernel<10, 2>, 10, 2);
            else if (ring_cfg == 112) e = launch_ring(dosage_moments_mma_ring_kernel<11, 2>, 11, 2);
            else if (ring_cfg == 63) e = launch_ring(dosage_moments_mma_ring_kernel<6, 3>, 6, 3);
            else if (ring_cfg == 73) e = launch_ring(dosage_moments_mma_ring_kernel<7, 3>, 7, 3);
            else {
            // attribute + occupancy query once per (device, shared-memory size): they are driver calls, not free per step
            static thread_local int cached_dev = -1, cached_occ = 1;
            static thread_local uint32_t cached_smem = 0;
            if (dev_now != cached_dev || pm.smem_bytes != cached_smem) {
                e = cudaFuncSetAttribute(dosage_moments_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         int(pm.smem_bytes));
                if (e != cudaSuccess) return e;
                int o = 1;
                if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, dosage_moments_mma_kernel, kMmaWarps * 32,
                                                                  pm.smem_bytes) != cudaSuccess || o < 1)
                    o = 1;
                cached_dev = dev_now;
                cached_smem = pm.smem_bytes;
                cached_occ = o;
            }
            int occ = cached_occ;
            const int occ_env = env_int("STOAT_GPU_K3M_CTAS", 0);
            if (occ_env >= 1 && occ_env < occ) occ = occ_env;
            uint64_t ctas = uint64_t(num_sms) * uint64_t(occ) / pm.n_slices;
            if (ctas < 1) ctas = 1;
            if (ctas > (n_tiles + kMmaWarps - 1) / kMmaWarps) ctas = (n_tiles + kMmaWarps - 1) / kMmaWarps;
            dosage_moments_mma_kernel<<<dim3(unsigned(ctas), pm.n_slices), kMmaWarps * 32, pm.smem_bytes, st>>>(q_main);
            e = cudaGetLastError();
            }
            if (e != cudaSuccess) return e;
            if ((e = cudaGetLastError()) != cudaSuccess) return e;
            if (ev_table_done && (e = cudaEventRecord(ev_table_done, st)) != cudaSuccess) return e;
            if (launches) *launches += 1;
            if (phase == 1) return cudaSuccess;
            if (three) {
                static thread_local int dev3 = -1, occ3 = 1;
                static thread_local uint32_t smem3 = 0;
                if (dev_now != dev3 || smem_small != smem3) {
                    e = cudaFuncSetAttribute(dosage_moments_mma3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             int(smem_small));
                    if (e != cudaSuccess) return e;
                    int o = 1;
                    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, dosage_moments_mma3_kernel, kMmaWarps * 32,
                                                                      smem_small) != cudaSuccess || o < 1)
                        o = 1;
                    dev3 = dev_now;
                    smem3 = smem_small;
                    occ3 = o;
                }
                const uint64_t tiles3 = (b.n_snarls + kMma3Tile - 1) / kMma3Tile;
                uint64_t ctas3 = uint64_t(num_sms) * uint64_t(occ3);
                if (ctas3 > (tiles3 + kMmaWarps - 1) / kMmaWarps) ctas3 = (tiles3 + kMmaWarps - 1) / kMmaWarps;
                dosage_moments_mma3_kernel<<<unsigned(ctas3), kMmaWarps * 32, smem_small, st>>>(q, rec3, b.counters + 5);
                if ((e = cudaGetLastError()) != cudaSuccess) return e;
                if (launches) *launches += 1;
            }
            if (thirds) {
                static thread_local int dev2 = -1, occ2 = 1;
                static thread_local uint32_t smem2 = 0;
                if (dev_now != dev2 || smem_small != smem2) {
                    e = cudaFuncSetAttribute(dosage_moments_mma2x_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             int(smem_small));
                    if (e != cudaSuccess) return e;
                    int o = 1;
                    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, dosage_moments_mma2x_kernel, kMmaWarps * 32,
                                                                      smem_small) != cudaSuccess || o < 1)
                        o = 1;
                    dev2 = dev_now;
                    smem2 = smem_small;
                    occ2 = o;
                }
                const uint64_t tiles2 = (b.n_snarls + kMma3Tile - 1) / kMma3Tile;
                uint64_t ctas2 = uint64_t(num_sms) * uint64_t(occ2);
                if (ctas2 > (tiles2 + kMmaWarps - 1) / kMmaWarps) ctas2 = (tiles2 + kMmaWarps - 1) / kMmaWarps;
                dosage_moments_mma2x_kernel<<<unsigned(ctas2), kMmaWarps * 32, smem_small, st>>>(q);
                if ((e = cudaGetLastError()) != cudaSuccess) return e;
                if (launches) *launches += 1;
            }
            p.list = slow_list;
            p.list_count = b.counters + 3;
            if (k3_ok)
                dosage_moments_tma_kernel<<<unsigned(num_sms), cfg.n_warps * 32, cfg.smem_bytes, st>>>(p);
            else
                slow_list_to_general_kernel<<<unsigned(num_sms), 256, 0, st>>>(slow_list, b.counters + 3, mom, b.big_list,
                                                                              b.counters);
            if ((e = cudaGetLastError()) != cudaSuccess) return e;
            if (launches) *launches += 1;
        } else if (k3_ok && part_i && part_d && slow_list && k3r_plan(pitch, cfg.stage_rows, pl)) {
            // K3R over every snarl (two-path snarls -> partial sums), then K3 over what it left on the slow list
            K3RParams q{};
            q.k = p;
            q.slice_words = pl.n_slices == 1 ? pitch / 4 : pl.slice_words;
            q.n_ring = pl.n_ring;
            q.slow_list = slow_list;
            q.slow_count = b.counters + 3;
            q.part = part;
            q.partial_only = 1;
            q.zero4[0] = q.zero4[1] = q.zero4[2] = q.zero4[3] = 0u;
            n_slices = pl.n_slices;
            if ((e = cudaMemsetAsync(b.counters + 3, 0, 4, st)) != cudaSuccess) return e;
            uint64_t ctas = uint64_t(num_sms) * pl.ctas_per_sm / pl.n_slices;
            if (ctas < 1) ctas = 1;
            if (ctas > b.n_snarls) ctas = b.n_snarls;
            k3r_kernel(pl.wpt)<<<dim3(unsigned(ctas), pl.n_slices), pl.threads, pl.smem_bytes, st>>>(q);
            if ((e = cudaGetLastError()) != cudaSuccess) return e;
            if (ev_table_done && (e = cudaEventRecord(ev_table_done, st)) != cudaSuccess) return e;
            p.list = slow_list;
            p.list_count = b.counters + 3;
            dosage_moments_tma_kernel<<<unsigned(num_sms), cfg.n_warps * 32, cfg.smem_bytes, st>>>(p);
            if ((e = cudaGetLastError()) != cudaSuccess) return e;
            if (launches) *launches += 1;
        } else {
            const uint64_t want = (b.n_snarls + cfg.n_warps - 1) / cfg.n_warps;
            const unsigned grid = unsigned(want < uint64_t(num_sms) ? want : uint64_t(num_sms));
            dosage_moments_tma_kernel<<<grid, cfg.n_warps * 32, cfg.smem_bytes, st>>>(p);
            if ((e = cudaGetLastError()) != cudaSuccess) return e;
            if (ev_table_done && (e = cudaEventRecord(ev_table_done, st)) != cudaSuccess) return e;
        }
        ols_solve_kernel<<<unsigned((b.n_snarls + kSolveChunk - 1) / kSolveChunk), kSolveThreads, 0, st>>>(
            mom, b.n_snarls, p, part, n_slices, b.snarl_path_off, out, b.big_list, b.counters, 0u, tail_t, tail_x, tail_df);
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
        if ((e = launch_student_tail(tail_t, tail_x, tail_df, b, out.p, num_sms, st)) != cudaSuccess) return e;
        if (launches) *launches += 1;
        g.list = b.big_list;
        g.n_list_dev = b.counters;
        if ((e = launch_quant_general(g, b.n_snarls, num_sms, st)) != cudaSuccess) return e;
        if (launches) *launches += 3;
        return cudaSuccess;
    }
    if (phase != 0) return cudaErrorNotSupported;  // covariates / K3M unavailable: no sample-sharded moments path
    if (variant == 0 && n_covar && cov_mom && cov_prefix && cov_rows >= n_samples && yc && tail_t && quant_covar_fast(n_covar)) {
        // covariates: K3C (one sweep per snarl) + K4C (solve) + K4b (tails), the general kernel on what they hand back.
        // The list is this path's own: it starts empty (prep's entries are shapes K3C turns away itself).
        CovParams c{};
        c.rows = m.rows;
        c.pitch = pitch;
        c.n_hap = uint32_t(m.n_cols);
        c.n_samples = n_samples;
        c.n_covar = n_covar;
        c.yc = yc;
        c.cov_rows = cov_rows;
        c.covar_t = cov_prefix + size_t(cov_rows + 1) * (n_covar + n_covar * (n_covar + 1) / 2);  // behind the prefix table
        c.prefix = cov_prefix;
        c.b = b;
        c.thr = thr;
        c.out = out;
        c.mom = static_cast<CovMoments*>(cov_mom);
        c.g_list = b.big_list;
        c.g_count = b.counters;
        if ((e = cudaMemsetAsync(b.counters, 0, 4, st)) != cudaSuccess) return e;
        if ((e = launch_covar_moments(c, num_sms, st)) != cudaSuccess) return e;
        if (ev_table_done && (e = cudaEventRecord(ev_table_done, st)) != cudaSuccess) return e;
        const SolveArgs sa{out, b.big_list, b.counters, 0u, tail_t, tail_x, tail_df, b.n_snarls};
        covar_solve_kernel<<<unsigned((b.n_snarls + 127) / 128), 128, 0, st>>>(c, sa);
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
        if ((e = launch_student_tail(tail_t, tail_x, tail_df, b, out.p, num_sms, st)) != cudaSuccess) return e;
        g.list = b.big_list;
        g.n_list_dev = b.counters;
        if ((e = launch_quant_general(g, b.n_snarls, num_sms, st)) != cudaSuccess) return e;
        // K3C instances (2-, 3-, 4-path snarls while 1 + D + Q <= 8), K4C, K4b, the general kernel's two launches
        if (launches) *launches += 1 + (3 + n_covar <= uint32_t(kCovFeat)) + (4 + n_covar <= uint32_t(kCovFeat)) + 1 + 1 + 2;
        return cudaSuccess;
    }
    g.list = nullptr;
    g.n_list_dev = nullptr;
    g.n_list_host = b.n_snarls;
    if (launches) *launches += 1;
    e = launch_quant_general(g, b.n_snarls, num_sms, st);
    if (e == cudaSuccess && ev_table_done) e = cudaEventRecord(ev_table_done, st);
    return e;
}

// centre + permute the phenotype:  yc[i] = y[i] - mean;  ys[(i % 16) * n_words + i / 16] = yc[i] (0 beyond n)
__global__ void centre_permute_kernel(const double* __restrict__ y, uint32_t n, double mean, uint32_t n_words,
                                      double* __restrict__ yc, double* __restrict__ ys) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= 16u * n_words) return;
    const double v = (i < n) ? y[i] - mean : 0.0;
    if (i < n) yc[i] = v;
    ys[(i % 16u) * n_words + i / 16u] = v;
}

cudaError_t launch_centre_permute(const double* y, uint32_t n, double mean, uint32_t n_words, double* yc, double* ys,
                                  cudaStream_t st) {
    const uint32_t total = 16u * n_words;
    if (total == 0) return cudaSuccess;
    centre_permute_kernel<<<(total + 255) / 256, 256, 0, st>>>(y, n, mean, n_words, yc, ys);
    return cudaGetLastError();
}

size_t quant_moments_bytes(uint64_t n_snarls) { return size_t(n_snarls) * sizeof(QuantMoments); }

// =================================================================================================
// eQTL: EQTLSnarlAnalyzer::analyze_and_write_snarl (src/snarl_analyzer.cpp:493-539) on the general path —
// snarls [s_begin, s_begin + n_snarls), per-pair outputs written at (pair index - item_base)
// =================================================================================================
cudaError_t launch_eqtl_general(const MatrixDev& m, const double* expression, uint32_t n_genes,
                                const uint64_t* snarl_gene_off, const uint32_t* gene_idx, const double* covar,
                                uint32_t n_covar, const BatchDev& b, uint64_t s_begin, uint64_t n_snarls,
                                uint64_t item_base, const stoat_gpu_thresholds& thr, QuantOut out, int num_sms,
                                cudaStream_t st, const uint32_t* list, const double* cov_prefix, uint32_t cov_rows,
                                uint8_t* cov_flag, uint64_t max_items_per_snarl) {
    if (n_snarls == 0) return cudaSuccess;
    // covariates: the sweep kernels of quant_covar.cuh first; they flag the entries the general kernel must take
    // (without covariates the same kernels run with Q = 0: the gene windows of eqtl_batch and the shapes a dense scan
    // leaves over; the dense scan itself is the tensor-core path's)
    const bool cov_fast = cov_flag && quant_covar_fast(n_covar ? n_covar : 1u) &&
                          (n_covar == 0 || (covar && cov_prefix && cov_rows >= uint32_t(m.n_cols / 2)));
    if (cov_fast) {
        CovEqtlParams q{};
        q.c.rows = m.rows;
        q.c.pitch = uint32_t(m.pitch);
        q.c.n_hap = uint32_t(m.n_cols);
        q.c.n_samples = uint32_t(m.n_cols / 2);
        q.c.n_covar = n_covar;
        q.c.cov_rows = cov_rows;
        q.c.covar_t = n_covar ? cov_prefix + size_t(cov_rows + 1) * (n_covar + n_covar * (n_covar + 1) / 2) : nullptr;
        q.c.b = b;
        q.c.thr = thr;
        q.c.out = out;
        q.prefix = cov_prefix;
        q.expression = expression;
        q.n_genes = n_genes;
        q.snarl_gene_off = snarl_gene_off;
        q.gene_idx = gene_idx;
        q.list = list;
        q.s_begin = s_begin;
        q.n_entries = n_snarls;
        q.item_base = item_base;
        q.flag = cov_flag;
        cudaError_t e = cudaMemsetAsync(cov_flag, 0, n_snarls, st);
        if (e != cudaSuccess) return e;
        if ((e = launch_eqtl_covar(q, max_items_per_snarl ? max_items_per_snarl : n_genes, num_sms, st)) != cudaSuccess) return e;
    }
    GParams g{};
    g.rows = m.rows;
    g.pitch = uint32_t(m.pitch);
    g.n_hap = uint32_t(m.n_cols);
    g.n_samples = uint32_t(m.n_cols / 2);
    g.y = expression;
    g.covar = covar;
    g.n_covar = n_covar;
    g.b = b;
    g.thr = thr;
    g.out = out;
    g.n_genes = n_genes;
    g.snarl_gene_off = snarl_gene_off;
    g.gene_idx = gene_idx;
    g.list = list;  // list != nullptr: snarls list[0 .. n_snarls), items of entry i at [i * n_genes, (i + 1) * n_genes)
    g.n_list_dev = nullptr;
    g.n_list_host = n_snarls;
    g.s_begin = s_begin;
    g.item_base = item_base;
    g.items_by_slot = list != nullptr;
    g.gene_split = 1;
    g.only_flagged = cov_fast ? cov_flag : nullptr;
    if (list) {  // few snarls, many genes each: give every snarl enough warps to fill the machine
        const uint64_t target = uint64_t(num_sms) * 3 * kGWarps * 2;
        const uint64_t want = (target + n_snarls - 1) / n_snarls, most = (uint64_t(n_genes) + 7) / 8;
        g.gene_split = uint32_t(std::max<uint64_t>(1, std::min(want, most)));
    }
    return launch_quant_general(g, n_snarls * g.gene_split, num_sms, st);
}

// BinaryCovarSnarlAnalyzer (src/snarl_analyzer.cpp:392-428) on the general path: the table against the binary phenotype,
// then the logistic regression of quant_general.cuh.  covar == nullptr reproduces the reference (covariates ignored).
cudaError_t launch_logit_general(const MatrixDev& m, const uint8_t* is_case, const double* covar, uint32_t n_covar,
                                 const BatchDev& b, const stoat_gpu_thresholds& thr, LogitOut out, const uint32_t* list,
                                 const uint32_t* n_list_dev, uint64_t n_max, int num_sms, cudaStream_t st) {
    if (n_max == 0) return cudaSuccess;
    GParams g{};
    g.rows = m.rows;
    g.pitch = uint32_t(m.pitch);
    g.n_hap = uint32_t(m.n_cols);
    g.n_samples = uint32_t(m.n_cols / 2);
    g.y = nullptr;
    g.is_case = is_case;
    g.covar = covar;
    g.n_covar = covar ? n_covar : 0;
    g.b = b;
    g.thr = thr;
    g.out = QuantOut{out.allele_paths, out.filtered, out.n_used, out.p, nullptr, out.beta, out.se, out.code};
    g.list = list;
    g.n_list_dev = list ? n_list_dev : nullptr;
    g.n_list_host = n_max;
    g.gene_split = 1;
    return launch_quant_general(g, n_max, num_sms, st);
}

// per-snarl reduction of a dense chunk: smallest p, its gene, number of genes with p < p_threshold
// (the device-side stand-in for writing 4e9 TSV rows; mirrors --table-threshold, src/snarl_analyzer.cpp:524-527)
__global__ void __launch_bounds__(128) eqtl_summary_kernel(const double* __restrict__ p, const uint8_t* __restrict__ code,
                                                           const uint8_t* __restrict__ filtered, uint64_t s_begin,
                                                           uint64_t n_snarls, uint32_t n_genes, double p_threshold,
                                                           double* __restrict__ min_p, uint32_t* __restrict__ argmin,
                                                           uint32_t* __restrict__ n_below, const uint32_t* __restrict__ list) {
    const uint32_t lane = threadIdx.x & 31;
    const uint64_t w = (uint64_t(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    if (w >= n_snarls) return;
    const uint64_t s = list ? uint64_t(list[w]) : s_begin + w;  // pair table row w <-> snarl s
    double best = __longlong_as_double(0x7ff0000000000000LL);  // +inf
    uint32_t best_g = 0xFFFFFFFFu, cnt = 0;
    if (!filtered[s]) {
        for (uint32_t gidx = lane; gidx < n_genes; gidx += 32) {
            const uint64_t k = w * uint64_t(n_genes) + gidx;
            if (code[k] != kPValue) continue;
            const double v = p[k];
            if (v < p_threshold) ++cnt;
            if (v < best || (v == best && gidx < best_g)) {
                best = v;
                best_g = gidx;
            }
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const double ob = __shfl_xor_sync(0xffffffffu, best, o);
        const uint32_t og = __shfl_xor_sync(0xffffffffu, best_g, o);
        if (ob < best || (ob == best && og < best_g)) {
            best = ob;
            best_g = og;
        }
    }
    cnt = __reduce_add_sync(0xffffffffu, cnt);
    if (lane == 0) {
        const bool none = best_g == 0xFFFFFFFFu;
        min_p[s] = none ? __longlong_as_double(0x7ff8000000000000LL) : best;
        argmin[s] = none ? 0u : best_g;
        n_below[s] = cnt;
    }
}

cudaError_t launch_eqtl_summary(const double* p, const uint8_t* code, const uint8_t* filtered, uint64_t s_begin,
                                uint64_t n_snarls, uint32_t n_genes, double p_threshold, double* min_p,
                                uint32_t* argmin, uint32_t* n_below, cudaStream_t st, const uint32_t* list) {
    if (n_snarls == 0) return cudaSuccess;
    const unsigned blocks = unsigned((n_snarls + 3) / 4);
    eqtl_summary_kernel<<<blocks, 128, 0, st>>>(p, code, filtered, s_begin, n_snarls, n_genes, p_threshold, min_p,
                                                argmin, n_below, list);
    return cudaGetLastError();
}

}  // namespace stoat
