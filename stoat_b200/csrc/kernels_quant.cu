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
#include "ctx.hpp"
#include "ols.cuh"
#include "ptx.cuh"

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
};

__device__ __forceinline__ uint32_t valid_word(uint32_t w, uint32_t n_hap) {
    const uint32_t lo = w * 32;
    if (lo + 32 <= n_hap) return 0xFFFFFFFFu;
    if (lo >= n_hap) return 0u;
    return (1u << (n_hap - lo)) - 1u;
}

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

// =================================================================================================
// K3 kernel: per-warp TMA pipelines (same structure as K1), fast path per snarl
// =================================================================================================
__global__ void __launch_bounds__(384, 1) dosage_moments_tma_kernel(const K3Params p) {
    extern __shared__ __align__(128) uint8_t smem[];
    const uint32_t n_warps = blockDim.x >> 5;
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
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

    const uint64_t gw = uint64_t(blockIdx.x) * n_warps + warp;
    const uint64_t stride = uint64_t(gridDim.x) * n_warps;
    uint64_t* my_full = full + warp * p.n_stages;
    uint8_t* my_stages = stages + size_t(warp) * p.n_stages * stage_bytes;
    const uint64_t policy = make_evict_first_policy();

    auto issue = [&](const SnarlRec& r, uint32_t stage) {
        uint8_t* dst = my_stages + size_t(stage) * stage_bytes;
        const uint32_t blob_bytes = uint32_t(r.blob_len16) * 16;
        mbar_arrive_expect_tx(&my_full[stage], blob_bytes + uint32_t(r.n_rows) * p.pitch);
        tma_bulk_g2s(dst, p.blob + size_t(r.blob16) * 16, blob_bytes, &my_full[stage]);
        for (uint32_t k = 0; k < r.n_rows; ++k)
            tma_bulk_g2s_hint(dst + kBlobBytesQ + k * p.pitch, p.rows + size_t(r.rows[k]) * p.pitch, p.pitch,
                              &my_full[stage], policy);
    };
    if (lane == 0) {
        uint64_t s = gw;
        for (uint32_t st = 0; st < p.n_stages && s < p.n_snarls; ++st, s += stride) issue(p.rec[s], st);
    }
    __syncwarp();

    uint32_t it = 0;
    for (uint64_t s = gw; s < p.n_snarls; s += stride, ++it) {
        const uint32_t stage = it % p.n_stages;
        const uint32_t parity = (it / p.n_stages) & 1u;
        const uint64_t s_next = s + uint64_t(p.n_stages) * stride;
        SnarlRec next_rec;
        if (lane == 0 && s_next < p.n_snarls) next_rec = p.rec[s_next];
        mbar_wait(&my_full[stage], parity);
        const uint8_t* blob = my_stages + size_t(stage) * stage_bytes;
        const uint8_t* rows_s = blob + kBlobBytesQ;
        const uint32_t np = blob[0];
        if (np) {  // np == 0: prep already put the snarl on the general list
            bool done = false;
            if (np <= 2)
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
        if (lane == 0 && s_next < p.n_snarls) issue(next_rec, stage);  // WAR on a read-only stage: no proxy fence
    }
}

// =================================================================================================
// K4: solve
// =================================================================================================
__global__ void __launch_bounds__(128) ols_solve_kernel(const QuantMoments* __restrict__ mom, uint64_t n_snarls,
                                                        QuantOut out, uint32_t* g_list, uint32_t* g_count) {
    const uint64_t s = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (s >= n_snarls) return;
    const QuantMoments m = mom[s];
    if (m.state != 1) return;
    const int p = m.p;
    const double nan = __longlong_as_double(0x7ff8000000000000LL);
    if (p < 2) {  // no regression column left: the reference indexes past the end (stats_test.cpp:487-489)
        out.p[s] = out.r2[s] = out.beta[s] = out.se[s] = nan;
        out.code[s] = kPUndef;
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
        const uint32_t slot = atomicAdd(g_count, 1u);
        g_list[slot] = uint32_t(s);
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
        const uint32_t slot = atomicAdd(g_count, 1u);
        g_list[slot] = uint32_t(s);
        return;
    }
    const OlsOut o = ols_statistics(beta[1], inv(1, 1), rss, tss, m.n_used, p);
    out.p[s] = o.p;
    out.r2[s] = o.r2;
    out.beta[s] = o.beta;
    out.se[s] = o.se;
    out.code[s] = kPValue;
}

// =================================================================================================
// K3G: general path, one warp per snarl, straight from global memory
// =================================================================================================
struct GParams {
    const uint8_t* rows;
    uint32_t pitch, n_hap, n_samples;
    const double* y;      // RAW phenotype (not centred: the pseudo-inverse branch is not shift invariant)
    const double* covar;  // [n_samples][n_covar] row-major (indexed by FILTERED row position, see below)
    uint32_t n_covar;
    BatchDev b;
    const uint32_t* list;  // nullptr: every snarl
    const uint32_t* n_list_dev;
    uint64_t n_list_host;  // used when list == nullptr
    stoat_gpu_thresholds thr;
    QuantOut out;
};

constexpr int kGWarps = 4;
struct GScratch {
    double tile[32][kMaxFeat + 1];   // X rows of the current 32 samples, last column = y
    double beta[kMaxFeat];
    double colsum[kMaxPathsG];
    uint8_t cnt[32][kMaxPathsG];     // haplotype counts of the current 32 samples per kept column
    uint8_t rs[32];                  // row sums
    uint8_t kept[kMaxPathsG];        // kept column -> path
    uint8_t fcol[kMaxPathsG];        // final column -> kept column
    uint8_t fmult[kMaxPathsG];
    uint8_t merged[kMaxPathsG];
    uint8_t same[kMaxPathsG];
};

// word `w` (32 haplotypes) of the presence vector of path `pth`
__device__ __forceinline__ uint32_t g_path_word(const GParams& g, uint64_t pth, uint32_t w) {
    const uint64_t a = g.b.path_edge_off[pth], z = g.b.path_edge_off[pth + 1];
    if (z == a) return valid_word(w, g.n_hap);
    uint32_t v = 0xFFFFFFFFu;
    for (uint64_t e = a; e < z; ++e) {
        const uint32_t r = g.b.edge_row[e];
        if (r == kMissingQ) return 0u;
        v &= __ldg(reinterpret_cast<const uint32_t*>(g.rows + size_t(r) * g.pitch) + w);
    }
    return v;
}

__device__ void quant_general(const GParams& g, GScratch& sc, uint64_t s, uint32_t lane) {
    const uint64_t p0 = g.b.snarl_path_off[s], p1 = g.b.snarl_path_off[s + 1];
    const uint32_t np = uint32_t(p1 - p0);
    const double nan = __longlong_as_double(0x7ff8000000000000LL);
    const uint32_t n_words = g.pitch / 4;
    auto finish_none = [&](bool filtered, uint32_t n_used, uint8_t code) {
        if (lane == 0) {
            g.out.filtered[s] = filtered ? 1 : 0;
            g.out.n_used[s] = n_used;
            g.out.p[s] = g.out.r2[s] = g.out.beta[s] = g.out.se[s] = nan;
            g.out.code[s] = code;
        }
    };
    if (np > uint32_t(kMaxPathsG)) {  // over the device limit: reported, never mangled
        for (uint32_t j = lane; j < np; j += 32) g.out.allele_paths[p0 + j] = 0;
        finish_none(true, 0, kPOverLimit);
        return;
    }
    // ---- allele_paths and the kept columns (quantitative_table.cpp:37-58) --------------------------
    uint32_t nk = 0;
    for (uint32_t j = 0; j < np; ++j) {
        uint32_t c = 0;
        for (uint32_t w = lane; w < n_words; w += 32) c += __popc(g_path_word(g, p0 + j, w));
        c = __reduce_add_sync(0xffffffffu, c);
        if (lane == 0) g.out.allele_paths[p0 + j] = c;
        if (c > 0) {
            if (lane == 0) sc.kept[nk] = uint8_t(j);
            ++nk;
        }
    }
    __syncwarp();
    const uint32_t n_blocks = (g.n_samples + 31) / 32;
    // counts of the 32 samples of block `blk` for every kept column -> sc.cnt / sc.rs (lane = sample)
    auto load_block = [&](uint32_t blk) -> uint32_t {
        const uint32_t i = blk * 32 + lane;
        const uint32_t w = blk * 2 + (lane >> 4);
        const uint32_t sh = (lane & 15) * 2;
        uint32_t rs = 0;
        for (uint32_t c = 0; c < nk; ++c) {
            const uint32_t v = (w < n_words) ? g_path_word(g, p0 + sc.kept[c], w) : 0u;
            const uint32_t k = ((v >> sh) & 1u) + ((v >> (sh + 1)) & 1u);
            sc.cnt[lane][c] = uint8_t(i < g.n_samples ? k : 0);
            rs += (i < g.n_samples ? k : 0);
        }
        sc.rs[lane] = uint8_t(rs > 255 ? 255 : rs);
        return rs;
    };
    // ---- sweep A: used rows, column sums in the reference's accumulation order (:558-563) ----------
    uint32_t n_used = 0;
    bool nondyadic = false;
    double colsum0 = 0.0, colsum1 = 0.0;  // lane <-> column (two columns per lane when nk > 32)
    for (uint32_t blk = 0; blk < n_blocks; ++blk) {
        const uint32_t rs = load_block(blk);
        const uint32_t used_ballot = __ballot_sync(0xffffffffu, rs > 0);
        n_used += __popc(used_ballot);
        nondyadic |= __any_sync(0xffffffffu, rs >= 3 && rs != 4 && rs != 8 && rs != 16 && rs != 32 && rs != 64);
        __syncwarp();
        for (uint32_t i = 0; i < 32; ++i) {
            if (!((used_ballot >> i) & 1u)) continue;
            const double rsum = double(sc.rs[i]);
            if (lane < nk) {
                const uint32_t k = sc.cnt[i][lane];
                colsum0 += (k > 0) ? __ddiv_rn(double(k), rsum) : 0.0;  // quantitative_table.cpp:93
            }
            if (lane + 32 < nk) {
                const uint32_t k = sc.cnt[i][lane + 32];
                colsum1 += (k > 0) ? __ddiv_rn(double(k), rsum) : 0.0;
            }
        }
        __syncwarp();
    }
    // remove_empty_columns is a no-op here (empty paths were never added); filtration (:549)
    if (n_used == 0 || nk < 2 || uint64_t(n_used) < g.thr.min_individuals) {
        finish_none(true, n_used, kPNa);
        return;
    }
    if (lane < nk) sc.colsum[lane] = colsum0;
    if (lane + 32 < nk) sc.colsum[lane + 32] = colsum1;
    __syncwarp();
    double total = 0.0;
    if (!nondyadic) {
        for (uint32_t c = 0; c < nk; ++c) total += sc.colsum[c];  // exact: every term is a multiple of 2^-6
    } else {
        // thirds etc.: reproduce the reference's single running sum over (row, column) (:558-563)
        if (lane == 0) {
            double t = 0.0;
            for (uint32_t i = 0; i < g.n_samples; ++i) {
                const uint32_t w = i >> 4, sh = (i & 15) * 2;
                uint32_t rs = 0;
                for (uint32_t c = 0; c < nk; ++c) {
                    const uint32_t v = g_path_word(g, p0 + sc.kept[c], w);
                    rs += ((v >> sh) & 1u) + ((v >> (sh + 1)) & 1u);
                }
                if (rs == 0) continue;
                for (uint32_t c = 0; c < nk; ++c) {
                    const uint32_t v = g_path_word(g, p0 + sc.kept[c], w);
                    const uint32_t k = ((v >> sh) & 1u) + ((v >> (sh + 1)) & 1u);
                    t = __dadd_rn(t, (k > 0) ? __ddiv_rn(double(k), double(rs)) : 0.0);
                }
            }
            total = t;
        }
        total = __shfl_sync(0xffffffffu, total, 0);
    }
    bool filtered = total < double(g.thr.min_haplotypes);  // :565
    if (!filtered) {
        int above = 0;
        for (uint32_t c = 0; c < nk; ++c) {  // :570-576
            const double freq = __ddiv_rn(sc.colsum[c], total);
            const double maf = fmin(freq, __dsub_rn(1.0, freq));
            if (maf > g.thr.maf_threshold) ++above;
        }
        filtered = above < 2;
    }
    if (filtered) {
        finish_none(true, n_used, kPNa);
        return;
    }
    // ---- combine identical columns (:620-674) ------------------------------------------------------
    uint32_t nf = 0;
    if (lane == 0)
        for (uint32_t c = 0; c < nk; ++c) sc.merged[c] = 0;
    __syncwarp();
    for (uint32_t a = 0; a < nk; ++a) {
        if (sc.merged[a]) continue;
        uint32_t mult = 1;
        if (nk >= 3 && a + 1 < nk) {
            // columns b > a identical to a over every used row (lane <-> candidate column b)
            bool same0 = (a + 1 + lane < nk) && !sc.merged[a + 1 + lane];
            bool same1 = (a + 33 + lane < nk) && !sc.merged[a + 33 + lane];
            if (__any_sync(0xffffffffu, same0 || same1)) {
                for (uint32_t blk = 0; blk < n_blocks; ++blk) {
                    load_block(blk);
                    __syncwarp();
                    for (uint32_t i = 0; i < 32; ++i) {
                        const uint8_t ka = sc.cnt[i][a];
                        if (same0 && sc.cnt[i][a + 1 + lane] != ka) same0 = false;
                        if (same1 && sc.cnt[i][a + 33 + lane] != ka) same1 = false;
                    }
                    __syncwarp();
                }
            }
            if (a + 1 + lane < nk) sc.same[a + 1 + lane] = same0 ? 1 : 0;
            if (a + 33 + lane < nk) sc.same[a + 33 + lane] = same1 ? 1 : 0;
            __syncwarp();
            if (lane == 0)
                for (uint32_t b2 = a + 1; b2 < nk; ++b2)
                    if (sc.same[b2]) sc.merged[b2] = 1;
            __syncwarp();
            for (uint32_t b2 = a + 1; b2 < nk; ++b2)
                if (sc.same[b2]) ++mult;
        }
        if (lane == 0) {
            sc.fcol[nf] = uint8_t(a);
            sc.fmult[nf] = uint8_t(mult);
        }
        ++nf;
        __syncwarp();
    }
    const uint32_t k_reg = nf - 1;  // remove_last_columns (:676-684)
    const uint32_t p = 1 + k_reg + g.n_covar;
    if (k_reg == 0) {  // the reference indexes past the end here (stats_test.cpp:487-489)
        finish_none(false, n_used, kPUndef);
        return;
    }
    if (p > uint32_t(kMaxFeat)) {
        finish_none(false, n_used, kPOverLimit);
        return;
    }
    // ---- sweep B: X^T X, X^T y ---------------------------------------------------------------------
    // entries e of the packed lower triangle + X^T y + y'y are owned by lanes (e % 32)
    const uint32_t n_tri = p * (p + 1) / 2;
    const uint32_t n_ent = n_tri + p + 2;  // + y'y + sum y
    double accv[6] = {0, 0, 0, 0, 0, 0};    // ceil((136 + 16 + 2) / 32) = 5
    uint32_t rank_base = 0;
    auto build_tile = [&](uint32_t blk, uint32_t used_ballot) {
        // lane = sample: write its X row (zeros when unused)
        const bool used = (used_ballot >> lane) & 1u;
        const uint32_t i = blk * 32 + lane;
        const uint32_t rank = rank_base + __popc(used_ballot & ((1u << lane) - 1u));
        const double rsum = double(sc.rs[lane]);
        sc.tile[lane][0] = used ? 1.0 : 0.0;
        for (uint32_t a = 0; a < k_reg; ++a) {
            const uint32_t k = sc.cnt[lane][sc.fcol[a]];
            // merged columns are sums of identical columns (:653-658): mult * (k / row_sum)
            double d = (used && k > 0) ? __ddiv_rn(double(k), rsum) : 0.0;
            double v = d;
            for (uint32_t m = 1; m < sc.fmult[a]; ++m) v = __dadd_rn(v, d);
            sc.tile[lane][1 + a] = v;
        }
        for (uint32_t c = 0; c < g.n_covar; ++c)
            // covariate rows are taken by position in the FILTERED table (stats_test.cpp:450-452)
            sc.tile[lane][1 + k_reg + c] = used ? g.covar[size_t(rank) * g.n_covar + c] : 0.0;
        sc.tile[lane][kMaxFeat] = (used && i < g.n_samples) ? g.y[i] : 0.0;
    };
    // (row, col) of the entries this lane owns
    uint32_t er[6], ec[6];
#pragma unroll
    for (int q = 0; q < 6; ++q) {
        const uint32_t e = lane + 32u * q;
        uint32_t r = 0;
        while ((r + 1) * (r + 2) / 2 <= e) ++r;
        er[q] = r;
        ec[q] = e - r * (r + 1) / 2;
    }
    for (uint32_t blk = 0; blk < n_blocks; ++blk) {
        const uint32_t rs = load_block(blk);
        const uint32_t used_ballot = __ballot_sync(0xffffffffu, rs > 0);
        __syncwarp();
        build_tile(blk, used_ballot);
        rank_base += __popc(used_ballot);
        __syncwarp();
#pragma unroll
        for (int q = 0; q < 6; ++q) {
            const uint32_t e = lane + 32u * q;
            if (e >= n_ent) continue;
            double a2 = 0.0;
            if (e < n_tri) {
                for (uint32_t i = 0; i < 32; ++i) a2 = fma(sc.tile[i][er[q]], sc.tile[i][ec[q]], a2);
            } else if (e < n_tri + p) {
                const uint32_t r = e - n_tri;
                for (uint32_t i = 0; i < 32; ++i) a2 = fma(sc.tile[i][r], sc.tile[i][kMaxFeat], a2);
            } else if (e == n_tri + p) {
                for (uint32_t i = 0; i < 32; ++i) a2 = fma(sc.tile[i][kMaxFeat], sc.tile[i][kMaxFeat], a2);
            } else {
                for (uint32_t i = 0; i < 32; ++i) a2 += sc.tile[i][kMaxFeat];
            }
            accv[q] += a2;
        }
        __syncwarp();
    }
    // ---- solve on lane 0 ----------------------------------------------------------------------------
    // gather the entries into lane 0 through shared memory (reuse the tile)
    __syncwarp();
    double* flat = &sc.tile[0][0];
#pragma unroll
    for (int q = 0; q < 6; ++q)
        if (lane + 32u * q < n_ent) flat[lane + 32u * q] = accv[q];
    __syncwarp();
    double inv11 = 0.0, sum_y = 0.0;
    if (lane == 0) {
        SmallMat<kMaxFeat> X, inv;
        uint32_t e = 0;
        for (uint32_t r = 0; r < p; ++r)
            for (uint32_t c = 0; c <= r; ++c, ++e) {
                X(r, c) = flat[e];
                X(c, r) = flat[e];
            }
        double xty[kMaxFeat];
        for (uint32_t r = 0; r < p; ++r) xty[r] = flat[n_tri + r];
        sum_y = flat[n_tri + p + 1];
        xtx_inverse<kMaxFeat>(X, int(p), inv);
        for (uint32_t r = 0; r < p; ++r) {
            double b2 = 0.0;
            for (uint32_t c = 0; c < p; ++c) b2 += inv(r, c) * xty[c];
            sc.beta[r] = b2;
        }
        inv11 = inv(1, 1);
    }
    __syncwarp();
    sum_y = __shfl_sync(0xffffffffu, sum_y, 0);
    const double y_mean = sum_y / double(n_used);
    // ---- sweep C: residuals (stats_test.cpp:459-465) -----------------------------------------------
    double rss = 0.0, tss = 0.0;
    rank_base = 0;
    for (uint32_t blk = 0; blk < n_blocks; ++blk) {
        const uint32_t rs = load_block(blk);
        const uint32_t used_ballot = __ballot_sync(0xffffffffu, rs > 0);
        __syncwarp();
        build_tile(blk, used_ballot);
        rank_base += __popc(used_ballot);
        __syncwarp();
        if ((used_ballot >> lane) & 1u) {
            double pred = 0.0;
            for (uint32_t r = 0; r < p; ++r) pred = fma(sc.tile[lane][r], sc.beta[r], pred);
            const double yv = sc.tile[lane][kMaxFeat];
            const double res = yv - pred;
            rss = fma(res, res, rss);
            const double dm = yv - y_mean;
            tss = fma(dm, dm, tss);
        }
        __syncwarp();
    }
    rss = warp_sum(rss);
    tss = warp_sum(tss);
    if (lane == 0) {
        const OlsOut o = ols_statistics(sc.beta[1], inv11, rss, tss, n_used, int(p));
        g.out.filtered[s] = 0;
        g.out.n_used[s] = n_used;
        g.out.p[s] = o.p;
        g.out.r2[s] = o.r2;
        g.out.beta[s] = o.beta;
        g.out.se[s] = o.se;
        g.out.code[s] = kPValue;
    }
    __syncwarp();
}

__global__ void __launch_bounds__(kGWarps * 32) quant_general_kernel(const GParams g) {
    __shared__ GScratch scratch[kGWarps];
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint64_t gw = uint64_t(blockIdx.x) * kGWarps + warp;
    const uint64_t n_warps = uint64_t(gridDim.x) * kGWarps;
    const uint64_t n = g.list ? uint64_t(*g.n_list_dev) : g.n_list_host;
    for (uint64_t i = gw; i < n; i += n_warps) {
        const uint64_t s = g.list ? g.list[i] : i;
        quant_general(g, scratch[warp], s, lane);
    }
}

// =================================================================================================
// host side
// =================================================================================================
struct K3Config {
    uint32_t n_warps, n_stages, stage_rows, smem_bytes;
};
static bool choose_k3_config(uint32_t pitch, uint32_t stage_rows, K3Config& cfg) {
    const uint32_t budget = 200 * 1024;
    if (stage_rows == 0) stage_rows = 1;
    const uint32_t stage_bytes = kBlobBytesQ + stage_rows * pitch;
    const uint32_t ys_bytes = (16u * (pitch / 4) * 8u + 127) & ~127u;
    if (ys_bytes + 2 * stage_bytes + 1024 > budget) return false;
    const uint32_t total_stages = (budget - ys_bytes - 1024) / stage_bytes;
    uint32_t n_stages = total_stages >= 12 ? 3 : 2;
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

// scratch owned by the caller: mom [n_snarls], ys permuted, sums
cudaError_t launch_quantitative_impl(const MatrixDev& m, const double* y_raw, const double* ys_perm,
                                     double sy_all, double syy_all, const double* covar, uint32_t n_covar,
                                     const BatchDev& b, uint32_t max_stage_rows, const stoat_gpu_thresholds& thr,
                                     QuantOut out, QuantMoments* mom, int num_sms, int variant, cudaStream_t st,
                                     uint32_t* launches) {
    if (b.n_snarls == 0) return cudaSuccess;
    const uint32_t pitch = uint32_t(m.pitch);
    const uint32_t n_samples = uint32_t(m.n_cols / 2);
    K3Config cfg{};
    // counters[0] already holds the number of snarls prep sent to the general list
    const bool fast = (variant == 0) && (n_covar == 0) && choose_k3_config(pitch, max_stage_rows, cfg);
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
        e = cudaFuncSetAttribute(dosage_moments_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 int(cfg.smem_bytes));
        if (e != cudaSuccess) return e;
        const uint64_t want = (b.n_snarls + cfg.n_warps - 1) / cfg.n_warps;
        const unsigned grid = unsigned(want < uint64_t(num_sms) ? want : uint64_t(num_sms));
        dosage_moments_tma_kernel<<<grid, cfg.n_warps * 32, cfg.smem_bytes, st>>>(p);
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
        ols_solve_kernel<<<unsigned((b.n_snarls + 127) / 128), 128, 0, st>>>(mom, b.n_snarls, out, b.big_list,
                                                                            b.counters);
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
        g.list = b.big_list;
        g.n_list_dev = b.counters;
        const uint64_t want_g = (b.n_snarls + kGWarps - 1) / kGWarps;
        const unsigned grid_g = unsigned(want_g < uint64_t(num_sms) * 4 ? want_g : uint64_t(num_sms) * 4);
        quant_general_kernel<<<grid_g, kGWarps * 32, 0, st>>>(g);
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
        if (launches) *launches += 3;
        return cudaSuccess;
    }
    g.list = nullptr;
    g.n_list_dev = nullptr;
    g.n_list_host = b.n_snarls;
    const uint64_t want_g = (b.n_snarls + kGWarps - 1) / kGWarps;
    const unsigned grid_g = unsigned(want_g < uint64_t(num_sms) * 8 ? want_g : uint64_t(num_sms) * 8);
    quant_general_kernel<<<grid_g, kGWarps * 32, 0, st>>>(g);
    if (launches) *launches += 1;
    return cudaGetLastError();
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

}  // namespace stoat
