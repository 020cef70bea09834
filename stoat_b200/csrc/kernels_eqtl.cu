// kernels_eqtl.cu — the DENSE eQTL scan (every snarl against every gene) as an FP64 tensor-core contraction.
//
// Reference: EQTLSnarlAnalyzer::analyze_and_write_snarl (src/snarl_analyzer.cpp:493-539) builds the snarl's dosage table
// once and then runs LinearRegression::linear_regression (src/stats_test.cpp:424-506) per gene, rebuilding X and
// re-inverting X^T X every time.  For the dominant snarl shape — two paths, so ONE dosage column d after the reference
// drops the last one (snarl_analyzer.cpp:676-684), X = [1, d] — everything that depends on the gene is
//     X^T y = [ sum_used y_g ,  sum d * y_g ],   y^T y = sum_used y_g^2
// and  sum d * y_g  over all (snarl, gene) pairs is the matrix product  D [S x n] * Y^T [n x G]:
//   E0  eqtl_dosage_codes   per snarl: dosage class of every sample as a 2-bit code (0, 1/2, 1 — exact), the reference's
//                           filters, (X^T X)^-1 by the reference's LDLT rule, the few samples that are on no path
//   EC  eqtl_centre         Yc = Y - mean_g (OLS with an intercept is shift invariant; removes cancellation), totals
//   E1  eqtl_dense_dmma     D * Yc^T with mma.sync m8n8k4 FP64 (SASS DMMA.8x8x4; D expanded from the 2-bit codes in
//                           registers, Yc staged by TMA bulk copies through a 4-stage mbarrier ring), epilogue per pair:
//                           2 x 2 solve, RSS by the moment expansion (explicit residuals when the fit is near perfect),
//                           t statistic; reduced on the fly to per-snarl best gene / count below the threshold
//   E2  eqtl_dense_finish   merge the gene groups, p = 2 * T_df(|t|) of the best gene
// Snarls of any other shape (three or more kept paths, non-dyadic dosages, rank-deficient designs, > 8 unused samples,
// covariates) go to the general kernel (quant_general.cuh), which follows the reference's order of operations.
#include <cstdint>

#include "ctx.hpp"
#include "dosage_masks.cuh"
#include "eqtl_dense.hpp"
#include "eqtl_pair.cuh"
#include "ols.cuh"
#include "ptx.cuh"

namespace stoat {

// =================================================================================================
// EC: centre the expression matrix, pad rows to a multiple of 64 samples
// =================================================================================================
__global__ void __launch_bounds__(256) eqtl_centre_kernel(const double* __restrict__ y, uint32_t n, uint32_t n_pad,
                                                          double* __restrict__ yc, double* __restrict__ tot,
                                                          double* __restrict__ tot2) {
    __shared__ double red[8];
    __shared__ double mean_s;
    const uint64_t g = blockIdx.x;
    const double* row = y + g * n;
    double* out = yc + g * n_pad;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double s = 0.0;
    for (uint32_t i = threadIdx.x; i < n; i += 256) s += row[i];
    s = warp_sum(s);
    if (lane == 0) red[warp] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < 8; ++w) t += red[w];
        mean_s = t / double(n);
    }
    __syncthreads();
    const double mean = mean_s;
    double a = 0.0, b = 0.0;
    for (uint32_t i = threadIdx.x; i < n_pad; i += 256) {
        const double v = (i < n) ? row[i] - mean : 0.0;
        out[i] = v;
        a += v;
        b = fma(v, v, b);
    }
    a = warp_sum(a);
    b = warp_sum(b);
    __syncthreads();
    if (lane == 0) red[warp] = a;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < 8; ++w) t += red[w];
        tot[g] = t;
    }
    __syncthreads();
    if (lane == 0) red[warp] = b;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < 8; ++w) t += red[w];
        tot2[g] = t;
    }
}

cudaError_t launch_eqtl_centre(const double* y, uint32_t n_genes, uint32_t n, uint32_t n_pad, double* yc, double* tot,
                               double* tot2, cudaStream_t st) {
    if (n_genes == 0) return cudaSuccess;
    eqtl_centre_kernel<<<n_genes, 256, 0, st>>>(y, n, n_pad, yc, tot, tot2);
    return cudaGetLastError();
}

// =================================================================================================
// E0: dosage codes, filters and (X^T X)^-1 per snarl — one warp per snarl
// =================================================================================================
// presence bits of chunk c (128 haplotypes = 64 samples) of one path: identify_path, src/snarl_analyzer.cpp:315-356
__device__ __forceinline__ uint4 eq_path_chunk(const MatrixDev& m, const BatchDev& b, uint64_t pth, uint32_t c) {
    const uint64_t a = b.path_edge_off[pth], z = b.path_edge_off[pth + 1];
    const uint32_t n_hap = uint32_t(m.n_cols);
    uint4 v = make_uint4(valid_word(4 * c, n_hap), valid_word(4 * c + 1, n_hap), valid_word(4 * c + 2, n_hap),
                         valid_word(4 * c + 3, n_hap));  // no checked edge: every haplotype (:343-354)
    for (uint64_t k = a; k < z; ++k) {
        const uint32_t r = b.edge_row[k];
        if (r == 0xFFFFFFFFu) return make_uint4(0, 0, 0, 0);  // edge absent from the matrix: nobody (:334-336)
        v = and4(v, __ldg(reinterpret_cast<const uint4*>(m.rows + uint64_t(r) * m.pitch) + c));
    }
    return v;
}

__global__ void __launch_bounds__(128) eqtl_dosage_codes_kernel(const MatrixDev m, const BatchDev b, const stoat_gpu_thresholds thr,
                                                                const bool allow_dense, const EqtlDense e, QuantOut out) {
    __shared__ uint16_t unused_s[4][kEqMaxUnused];
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint64_t s = uint64_t(blockIdx.x) * 4 + warp;
    if (s >= e.n_snarls) return;
    auto to_general = [&]() {
        if (lane == 0) e.fb_list[atomicAdd(e.fb_count, 1u)] = uint32_t(s);
    };
    const uint64_t p0 = b.snarl_path_off[s], p1 = b.snarl_path_off[s + 1];
    if (!allow_dense || p1 - p0 != 2 || e.n_samples > 65535u) return to_general();
    const uint32_t chunks = uint32_t(m.pitch / 16), n_hap = uint32_t(m.n_cols);
    uint32_t al0 = 0, al1 = 0, n_used = 0, n_full0 = 0, n_half = 0, nd = 0, n_un = 0;
    for (uint32_t c0 = 0; c0 < chunks; c0 += 32) {
        const uint32_t c = c0 + lane;
        uint32_t my_un = 0, un_mask[4] = {0, 0, 0, 0};
        if (c < chunks) {
            const uint4 a = eq_path_chunk(m, b, p0, c), bb = eq_path_chunk(m, b, p0 + 1, c);
            al0 += popc4(a);
            al1 += popc4(bb);
            const uint32_t av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {bb.x, bb.y, bb.z, bb.w};
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const LeanMasks mk = lean_masks(av[k], bv[k]);
                n_used += __popc(mk.used);
                n_full0 += __popc(mk.full0);
                n_half += __popc(mk.half);
                nd |= mk.third | mk.tthird;
                un_mask[k] = ~mk.used & 0x55555555u & valid_word(4 * c + k, n_hap);
                my_un += __popc(un_mask[k]);
            }
        }
        // samples on no path: keep the first kEqMaxUnused (in sample order)
        uint32_t incl = my_un;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= uint32_t(o)) incl += t;
        }
        uint32_t pos = n_un + incl - my_un;
        for (int k = 0; k < 4 && my_un; ++k) {
            uint32_t w = un_mask[k];
            while (w) {
                const int bit = __ffs(w) - 1;
                w &= w - 1;
                if (pos < uint32_t(kEqMaxUnused)) unused_s[warp][pos] = uint16_t(64 * c + 16 * k + (bit >> 1));
                ++pos;
            }
        }
        n_un += __shfl_sync(0xffffffffu, incl, 31);
    }
    __syncwarp();
    al0 = __reduce_add_sync(0xffffffffu, al0);
    al1 = __reduce_add_sync(0xffffffffu, al1);
    n_used = __reduce_add_sync(0xffffffffu, n_used);
    n_full0 = __reduce_add_sync(0xffffffffu, n_full0);
    n_half = __reduce_add_sync(0xffffffffu, n_half);
    if (__reduce_or_sync(0xffffffffu, nd) != 0u) return to_general();  // a sample on three path-haplotypes: thirds
    // ---- filters (snarl_analyzer.cpp:542-579, via remove_empty_columns :581-618); as K3, quant_lean.cuh ------------
    const uint32_t n_full1 = n_used - n_full0 - n_half;
    const double colsum0 = double(n_full0) + 0.5 * double(n_half);
    const double colsum1 = double(n_full1) + 0.5 * double(n_half);
    bool filtered = (n_used == 0) || (al0 == 0) || (al1 == 0) || (uint64_t(n_used) < thr.min_individuals);
    if (!filtered) {
        const double total = double(n_used);  // every used row sums to 1
        if (total < double(thr.min_haplotypes)) filtered = true;
        const double f0 = __ddiv_rn(colsum0, total), f1 = __ddiv_rn(colsum1, total);
        const double maf0 = fmin(f0, __dsub_rn(1.0, f0)), maf1 = fmin(f1, __dsub_rn(1.0, f1));
        if (!(maf0 > thr.maf_threshold && maf1 > thr.maf_threshold)) filtered = true;
    }
    if (lane < 2) out.allele_paths[p0 + lane] = lane ? al1 : al0;  // quantitative_table.cpp:54-55
    if (lane == 0) {
        out.filtered[s] = filtered ? 1 : 0;
        out.n_used[s] = n_used;
    }
    if (filtered) {  // reported here: the scan never sees this snarl
        if (lane == 0) {
            e.min_p[s] = __longlong_as_double(0x7ff8000000000000LL);
            e.argmin[s] = 0;
            e.n_below[s] = 0;
        }
        return;
    }
    if (n_un > uint32_t(kEqMaxUnused)) return to_general();
    uint32_t slot = 0xFFFFFFFFu;
    if (lane == 0) {
        SmallMat<4> X, inv;
        for (int i = 0; i < 16; ++i) X.a[i] = 0.0;
        const double sdd = double(n_full0) + 0.25 * double(n_half);
        X(0, 0) = double(n_used);
        X(0, 1) = X(1, 0) = colsum0;
        X(1, 1) = sdd;
        if (xtx_inverse<4>(X, 2, inv)) {  // rank deficient: the pseudo-inverse is not shift invariant -> raw y, general path
            e.fb_list[atomicAdd(e.fb_count, 1u)] = uint32_t(s);
        } else {
            // dense snarls are compacted: the contraction runs over slots, not over filtered / general-path snarls
            slot = atomicAdd(e.dense_count, 1u);
            EqtlRec rec{};
            rec.inv00 = inv(0, 0);
            rec.inv01 = inv(0, 1);
            rec.inv11 = inv(1, 1);
            rec.sd = colsum0;
            rec.sdd = sdd;
            rec.n_used = n_used;
            rec.n_unused = uint8_t(n_un);
            for (uint32_t k = 0; k < n_un; ++k) rec.unused[k] = unused_s[warp][k];
            rec.state = 1;
            e.rec[slot] = rec;
            e.slot_snarl[slot] = uint32_t(s);
        }
    }
    slot = __shfl_sync(0xffffffffu, slot, 0);
    if (slot == 0xFFFFFFFFu) return;
    // second pass over the (cache-hot) rows: the 2-bit codes, written at the snarl's slot
    for (uint32_t c = lane; c < chunks; c += 32) {
        const uint4 a = eq_path_chunk(m, b, p0, c), bb = eq_path_chunk(m, b, p0 + 1, c);
        const uint32_t av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {bb.x, bb.y, bb.z, bb.w};
        uint32_t cw[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const LeanMasks mk = lean_masks(av[k], bv[k]);
            cw[k] = (mk.full0 << 1) | mk.half;  // sample i of the word at bits 2i..2i+1: 0 -> 0, 1 -> 1/2, 2 -> 1
        }
        *reinterpret_cast<uint4*>(e.dcode + (uint64_t(c) * e.s_pad + slot) * 16) = make_uint4(cw[0], cw[1], cw[2], cw[3]);
    }
}

cudaError_t launch_eqtl_dosage_codes(const MatrixDev& m, const BatchDev& b, const stoat_gpu_thresholds& thr, bool allow_dense,
                                     const EqtlDense& e, QuantOut out, cudaStream_t st) {
    if (e.n_snarls == 0) return cudaSuccess;
    eqtl_dosage_codes_kernel<<<unsigned((e.n_snarls + 3) / 4), 128, 0, st>>>(m, b, thr, allow_dense, e, out);
    return cudaGetLastError();
}

// =================================================================================================
// E1: D * Yc^T on the FP64 tensor cores + per-pair OLS epilogue, reduced per snarl
// =================================================================================================
constexpr int kEqStages = 4;
constexpr int kEqBRow = kEqKT * 8 + 64;               // 576 B: row stride = 144 words = 16 (mod 32) -> LDS.128 conflict free
constexpr int kEqStageB = kEqTileG * kEqBRow;         // 36,864
constexpr int kEqStageA = kEqTileS * 16;              // 2,048
constexpr int kEqStageBytes = kEqStageB + kEqStageA;  // 38,912
constexpr int kEqConsumerWarps = 8;                   // 4 (snarls) x 2 (genes), warp tile 32 x 32
constexpr int kEqThreads = (kEqConsumerWarps + 1) * 32;

struct EqSmem {
    uint64_t full[kEqStages], empty[kEqStages];
    EqConst k;
    EqtlRec rec[kEqTileS];
    double best_key[2][kEqTileS];
    uint32_t best_gene[2][kEqTileS];
    uint32_t cnt[2][kEqTileS];
};
constexpr size_t kEqStageOff = (sizeof(EqSmem) + 127) / 128 * 128;
size_t eqtl_dense_smem_bytes() { return kEqStageOff + size_t(kEqStages) * kEqStageBytes; }

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}
__device__ __forceinline__ void named_bar_sync(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }

__global__ void __launch_bounds__(kEqThreads, 1) eqtl_dense_dmma_kernel(const EqtlDense e) {
    extern __shared__ __align__(128) uint8_t eq_smem_raw[];
    EqSmem& sm = *reinterpret_cast<EqSmem*>(eq_smem_raw);
    uint8_t* stages = eq_smem_raw + kEqStageOff;  // plain offset from the shared array: the compiler emits LDS, not generic LD
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint64_t s0 = uint64_t(blockIdx.x) * kEqTileS;  // first SLOT of this tile (dense snarls, compacted by E0)
    const uint32_t n_dense = *e.dense_count;
    if (s0 >= n_dense) return;
    const uint32_t grp = blockIdx.y;
    const uint32_t g_begin = grp * e.genes_per_group;
    const uint32_t g_end = min(e.n_genes, g_begin + e.genes_per_group);
    const uint32_t n_gt = (g_end - g_begin + kEqTileG - 1) / kEqTileG;
    const uint32_t n_it = n_gt * e.n_kst;

    if (tid == 0) {
        for (int i = 0; i < kEqStages; ++i) {
            mbar_init(&sm.full[i], 1);
            mbar_init(&sm.empty[i], kEqConsumerWarps);
        }
        mbar_fence_init();
        sm.k = EqConst{e.yc, e.tot, e.tot2, e.dcode, e.s_pad, e.n_pad, e.n_samples, e.p_threshold, e.z_threshold};
    }
    for (uint32_t i = tid; i < kEqTileS * (sizeof(EqtlRec) / 16); i += kEqThreads) {
        const bool live = s0 + i / (sizeof(EqtlRec) / 16) < n_dense;  // slots past the end: state 0, never reported
        reinterpret_cast<uint4*>(sm.rec)[i] = live ? reinterpret_cast<const uint4*>(e.rec + s0)[i] : make_uint4(0, 0, 0, 0);
    }
    for (uint32_t i = tid; i < 2 * kEqTileS; i += kEqThreads) {
        (&sm.best_key[0][0])[i] = -1.0;
        (&sm.best_gene[0][0])[i] = 0xFFFFFFFFu;
        (&sm.cnt[0][0])[i] = 0u;
    }
    __syncthreads();

    if (warp == kEqConsumerWarps) {
        // ---------------- producer warp: TMA bulk copies, one stage = 64 samples of 64 genes + the tile's dosage codes
        for (uint32_t it = 0; it < n_it; ++it) {
            const uint32_t st = it % kEqStages, round = it / kEqStages;
            if (round) mbar_wait(&sm.empty[st], (round - 1) & 1);
            const uint32_t gt = it / e.n_kst, kst = it - gt * e.n_kst;
            uint8_t* dst = stages + size_t(st) * kEqStageBytes;
            if (lane == 0) mbar_arrive_expect_tx(&sm.full[st], kEqTileG * kEqKT * 8 + kEqStageA);
            __syncwarp();
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const uint32_t row = lane * 2 + j;
                const uint32_t gene = min(g_begin + gt * kEqTileG + row, e.n_genes - 1);  // past the end: a valid row, ignored
                tma_bulk_g2s(dst + row * kEqBRow, e.yc + uint64_t(gene) * e.n_pad + uint64_t(kst) * kEqKT, kEqKT * 8, &sm.full[st]);
            }
            if (lane == 0) tma_bulk_g2s(dst + kEqStageB, e.dcode + (uint64_t(kst) * e.s_pad + s0) * 16, kEqStageA, &sm.full[st]);
        }
        return;
    }

    // ---------------- consumer warps
    const uint32_t wm = warp >> 1, wn = warp & 1, gid = lane >> 2, tig = lane & 3;
    double acc[4][4][2];
    uint32_t it = 0;
    for (uint32_t gt = 0; gt < n_gt; ++gt) {
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;
        for (uint32_t kst = 0; kst < e.n_kst; ++kst, ++it) {
            const uint32_t st = it % kEqStages;
            mbar_wait(&sm.full[st], (it / kEqStages) & 1);
            const uint8_t* sb = stages + size_t(st) * kEqStageBytes;
            const uint8_t* sa = sb + kEqStageB;
            uint4 aw[4];
#pragma unroll
            for (int mt = 0; mt < 4; ++mt) aw[mt] = *reinterpret_cast<const uint4*>(sa + (wm * 32 + mt * 8 + gid) * 16);
            const uint8_t* brow = sb + (wn * 32 + gid) * kEqBRow + tig * 16;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                // this thread's k index of the step pair (2j, 2j+1) is sample 8j + 2 tig (+1): any fixed permutation of
                // the samples is a valid contraction order, and this one makes the B operand one LDS.128
                double2 bv[4];
#pragma unroll
                for (int nt = 0; nt < 4; ++nt) bv[nt] = *reinterpret_cast<const double2*>(brow + nt * 8 * kEqBRow + j * 64);
                const uint32_t sh = 2 * (8 * (j & 1) + 2 * tig);
                double a_even[4], a_odd[4];
#pragma unroll
                for (int mt = 0; mt < 4; ++mt) {
                    const uint32_t w = (j < 2) ? aw[mt].x : (j < 4) ? aw[mt].y : (j < 6) ? aw[mt].z : aw[mt].w;
                    a_even[mt] = code_to_double((w >> sh) & 3u);
                    a_odd[mt] = code_to_double((w >> (sh + 2)) & 3u);
                }
                // 16 independent accumulator tiles between two DMMAs on the same tile (no back-to-back dependency)
#pragma unroll
                for (int mt = 0; mt < 4; ++mt)
#pragma unroll
                    for (int nt = 0; nt < 4; ++nt) dmma884(acc[mt][nt][0], acc[mt][nt][1], a_even[mt], bv[nt].x);
#pragma unroll
                for (int mt = 0; mt < 4; ++mt)
#pragma unroll
                    for (int nt = 0; nt < 4; ++nt) dmma884(acc[mt][nt][0], acc[mt][nt][1], a_odd[mt], bv[nt].y);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&sm.empty[st]);
        }
        // ---------------- epilogue of this gene tile: acc[mt][nt][c] = sum d * yc of (snarl, gene)
        const uint32_t g_tile = g_begin + gt * kEqTileG + wn * 32;
#pragma unroll
        for (int mt = 0; mt < 4; ++mt) {
            const uint32_t sl = wm * 32 + mt * 8 + gid;
            const EqtlRec* r = &sm.rec[sl];
            double bk = -1.0;
            uint32_t bg = 0xFFFFFFFFu, cnt = 0;
            if (r->state == 1) {
#pragma unroll
                for (int nt = 0; nt < 4; ++nt)
#pragma unroll
                    for (int c = 0; c < 2; ++c) {
                        const uint32_t g = g_tile + nt * 8 + 2 * tig + c;
                        if (g < g_end) {
                            const double key = eq_pair_stat(&sm.k, r, s0 + sl, g, acc[mt][nt][c]);  // < 0: also below threshold
                            const double ak = fabs(key) - 1.0;
                            cnt += key < 0.0 ? 1u : 0u;
                            if (ak > bk || (ak == bk && g < bg)) {
                                bk = ak;
                                bg = g;
                            }
                        }
                    }
            }
            // the four lanes of a quad hold the 32 genes of this snarl row
#pragma unroll
            for (int o = 1; o <= 2; o <<= 1) {
                const double ok = __shfl_xor_sync(0xffffffffu, bk, o);
                const uint32_t og = __shfl_xor_sync(0xffffffffu, bg, o);
                cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
                if (ok > bk || (ok == bk && og < bg)) {
                    bk = ok;
                    bg = og;
                }
            }
            if (tig == 0 && r->state == 1) {  // this warp alone owns [wn][its 32 snarls]
                sm.cnt[wn][sl] += cnt;
                const double cur = sm.best_key[wn][sl];
                if (bk > cur || (bk == cur && bg < sm.best_gene[wn][sl])) {
                    sm.best_key[wn][sl] = bk;
                    sm.best_gene[wn][sl] = bg;
                }
            }
        }
    }
    named_bar_sync(1, kEqConsumerWarps * 32);
    if (tid < kEqTileS) {
        double k0 = sm.best_key[0][tid], k1 = sm.best_key[1][tid];
        uint32_t ga = sm.best_gene[0][tid], gb = sm.best_gene[1][tid];
        if (k1 > k0 || (k1 == k0 && gb < ga)) {
            k0 = k1;
            ga = gb;
        }
        const uint64_t o = uint64_t(grp) * e.s_pad + s0 + tid;
        e.part_key[o] = k0;
        e.part_gene[o] = ga;
        e.part_cnt[o] = sm.cnt[0][tid] + sm.cnt[1][tid];
    }
}

cudaError_t launch_eqtl_dense(const EqtlDense& e, int /*num_sms*/, cudaStream_t st) {
    if (e.s_pad == 0 || e.n_genes == 0) return cudaSuccess;
    const int smem = int(eqtl_dense_smem_bytes());
    cudaError_t err = cudaFuncSetAttribute(eqtl_dense_dmma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (err != cudaSuccess) return err;
    // x = snarl tiles (fastest), y = gene groups: every tile passes over one L2-resident group before the next starts
    dim3 grid(unsigned(e.s_pad / kEqTileS), e.n_groups);
    eqtl_dense_dmma_kernel<<<grid, kEqThreads, smem, st>>>(e);
    return cudaGetLastError();
}

// =================================================================================================
// E2: merge the gene groups, p of the best gene
// =================================================================================================
__global__ void __launch_bounds__(128) eqtl_dense_finish_kernel(const EqtlDense e) {
    const uint64_t slot = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (slot >= *e.dense_count) return;
    const EqtlRec r = e.rec[slot];
    const uint32_t s = e.slot_snarl[slot];
    const double nan = __longlong_as_double(0x7ff8000000000000LL);
    double bk = -1.0;
    uint32_t bg = 0xFFFFFFFFu, cnt = 0;
    for (uint32_t grp = 0; grp < e.n_groups; ++grp) {
        const uint64_t o = uint64_t(grp) * e.s_pad + slot;
        const double k = e.part_key[o];
        const uint32_t g = e.part_gene[o];
        cnt += e.part_cnt[o];
        if (k > bk || (k == bk && g < bg)) {
            bk = k;
            bg = g;
        }
    }
    int df_res = int(r.n_used) - 2 + 1;
    if (df_res < 1) df_res = 1;
    const bool none = bg == 0xFFFFFFFFu;
    e.min_p[s] = none ? nan : 2.0 * student_t_upper_tail(double(df_res), bk);
    e.argmin[s] = none ? 0u : bg;
    e.n_below[s] = cnt;
}

cudaError_t launch_eqtl_dense_finish(const EqtlDense& e, cudaStream_t st) {
    if (e.n_snarls == 0) return cudaSuccess;
    eqtl_dense_finish_kernel<<<unsigned((e.n_snarls + 127) / 128), 128, 0, st>>>(e);
    return cudaGetLastError();
}

}  // namespace stoat
