// kernels_eqtl_umma.cu — the dense eQTL contraction on the 5th-generation tensor cores (tcgen05.mma kind::i8, accumulators in
// tensor memory): SURVEY §7's exact integer formulation of  D [S x n] * Yc^T [n x G].
//
// Reference: EQTLSnarlAnalyzer::analyze_and_write_snarl (src/snarl_analyzer.cpp:493-539), one LinearRegression per (snarl,
// gene); for a two-path snarl everything gene-dependent is sum_i d_i * y_gi (kernels_eqtl.cu).  Here that sum is EXACT:
//   A  = 2 d in {0, 1, 2} as int8                                  [slots  x samples]   (from E0's 2-bit dosage codes)
//   B  = the 8 signed 7-bit limbs of the centred expression        [genes * 8 x samples]
//        yc_gi = scale_g * sum_t q_gti * 128^-(t+1),  |q| <= 64, residual <= 2^-57 scale_g   (as K3M slices the phenotype)
//   C_gt = sum_i A_si * q_gti   (INT32: |C| <= 2 * 64 * n)         =>  sum_i d_i yc_gi = scale_g / 2 * sum_t C_gt 128^-(t+1)
// so the result does not depend on the order of summation (bit-reproducible under any tiling / sharding) and carries 56
// bits, more than an FP64 dot product of 5,000 terms keeps.  The FP64 form (eqtl_dense_dmma, DMMA) is bound by the FP64
// tensor rate (31 of cuBLAS DGEMM's 35 TFLOP/s); the int8 form does 8 x the multiply-adds at > 60 x the rate.
//
//   EU0a eqtl_gene_scale     scale_g = 2^(ilogb(max |yc_g|) + 2)
//   EU0b eqtl_limbs          Yc -> B, written as the images of the kernel's shared-memory tiles (below)
//   EU0c eqtl_expand_codes   2-bit codes -> A, same
//   EU1  eqtl_dense_umma     CTA = 128 slots x one gene group; per gene tile of 32 genes (N = 256 limb columns):
//        warp 4   producer: per 128-sample block ONE bulk copy (cp.async.bulk, UBLKCP) of the 16 KB A image and one of the
//                 32 KB B image into a 4-stage ring (mbarrier complete_tx) — the operands are stored in global memory
//                 exactly as the tensor core wants them in shared memory (K-major, 128-byte swizzle), so no tensor map
//        warp 5   MMA issuer: 4 x tcgen05.mma.cta_group::1.kind::i8 (M 128, N 256, K 32) per stage from shared-memory
//                 descriptors, tcgen05.commit frees the stage; the accumulator is one of TWO 256-column TMEM buffers
//        warps 0-3 epilogue, thread = slot: tcgen05.ld of a gene's 8 limb columns, Horner to FP64, then the same per-pair
//                 OLS tail as the FP64 kernel (eq_pair_stat) and the per-slot best gene / count, while the MMA warp is
//                 already filling the other TMEM buffer with the next gene tile
// Results are the FP64 kernel's to ~1e-13 (it rounds inside its dot products, this one does not).
#include <cstdint>

#include "ctx.hpp"
#include "eqtl_dense.hpp"
#include "eqtl_pair.cuh"
#include "ols.cuh"
#include "ptx.cuh"

namespace stoat {

// =================================================================================================
// operand images
// =================================================================================================
// byte (row r, k) of a K-major tile with 128-byte rows under the 128-byte swizzle: 16-byte chunk index XOR (row mod 8)
__host__ __device__ __forceinline__ uint32_t umma_swz(uint32_t r, uint32_t k) {
    return r * 128u + ((((k >> 4) ^ (r & 7u)) << 4) | (k & 15u));
}

__global__ void __launch_bounds__(256) eqtl_gene_scale_kernel(const double* __restrict__ yc, uint32_t n_pad,
                                                              double* __restrict__ scale) {
    __shared__ double red[8];
    const double* row = yc + uint64_t(blockIdx.x) * n_pad;
    double m = 0.0;
    for (uint32_t i = threadIdx.x; i < n_pad; i += 256) m = fmax(m, fabs(row[i]));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 8; ++w) m = fmax(m, red[w]);
        int ex = 0;
        if (m > 0.0 && isfinite(m)) frexp(m, &ex);  // m = f * 2^ex, f in [0.5, 1)
        scale[blockIdx.x] = ldexp(1.0, ex + 1);      // |yc| / scale < 0.5
    }
}

// thread = (gene, group of 16 samples): the 8 limbs of 16 values -> eight 16-byte chunks of the B image
__global__ void __launch_bounds__(256) eqtl_limbs_kernel(const double* __restrict__ yc, const double* __restrict__ scale,
                                                         uint32_t n_genes, uint32_t n_pad, uint32_t n_kb,
                                                         uint8_t* __restrict__ b_img) {
    const uint64_t t = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    const uint32_t groups = n_pad / 16;
    const uint32_t g_pad = (n_genes + kUmTileG - 1) / kUmTileG * kUmTileG;
    if (t >= uint64_t(g_pad) * groups) return;
    const uint32_t g = uint32_t(t / groups), grp = uint32_t(t % groups);
    int8_t q[8][16];
    if (g < n_genes) {
        const double inv = 1.0 / scale[g];  // a power of two: exact
        const double* src = yc + uint64_t(g) * n_pad + grp * 16;
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            double r = src[i] * inv;
#pragma unroll
            for (int l = 0; l < 8; ++l) {
                r *= 128.0;
                const double qv = rint(r);
                q[l][i] = int8_t(int(qv));
                r -= qv;
            }
        }
    } else {
#pragma unroll
        for (int l = 0; l < 8; ++l)
#pragma unroll
            for (int i = 0; i < 16; ++i) q[l][i] = 0;
    }
    const uint32_t gt = g / kUmTileG, gl = g % kUmTileG, kb = (grp * 16) / kUmKB, k = (grp * 16) % kUmKB;
    uint8_t* tile = b_img + (uint64_t(gt) * n_kb + kb) * kUmBBytes;
#pragma unroll
    for (int l = 0; l < 8; ++l) {
        uint4 w;
        memcpy(&w, q[l], 16);
        *reinterpret_cast<uint4*>(tile + umma_swz(gl * 8 + l, k)) = w;
    }
}

// thread = (slot, word of 16 samples): 2-bit codes (0, 1, 2 = 2 d) -> sixteen int8 of the A image
__global__ void __launch_bounds__(256) eqtl_expand_codes_kernel(const uint8_t* __restrict__ dcode, uint64_t s_pad,
                                                                uint32_t n_kst, uint32_t n_kb,
                                                                const uint32_t* __restrict__ dense_count,
                                                                uint8_t* __restrict__ a_img) {
    const uint64_t t = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    const uint32_t words = n_kst * 4;
    const uint64_t slot = t / words;
    const uint32_t w = uint32_t(t % words);
    // slots past the dense count hold stale codes: zero them (their rows are never reported, but keep the image defined)
    const uint64_t n_tiles = (uint64_t(*dense_count) + kUmTileS - 1) / kUmTileS;
    if (slot >= n_tiles * kUmTileS || slot >= s_pad) return;
    const uint32_t kst = w / 4, j = w % 4;
    uint32_t code = *reinterpret_cast<const uint32_t*>(dcode + (uint64_t(kst) * s_pad + slot) * 16 + j * 4);
    if (slot >= *dense_count) code = 0;
    uint8_t b[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) b[i] = uint8_t((code >> (2 * i)) & 3u);
    const uint32_t sample = kst * 64 + j * 16;
    const uint32_t kb = sample / kUmKB, k = sample % kUmKB;
    uint8_t* tile = a_img + ((slot / kUmTileS) * n_kb + kb) * kUmABytes;
    uint4 v;
    memcpy(&v, b, 16);
    *reinterpret_cast<uint4*>(tile + umma_swz(uint32_t(slot % kUmTileS), k)) = v;
}

// =================================================================================================
// tcgen05 / TMEM wrappers (PTX ISA 8.6+, sm_100a)
// =================================================================================================
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void named_bar_sync_um(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }
__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem, uint32_t cols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t cols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols) : "memory");
}
// D[tmem] (+)= A[smem] * B[smem], int8 x int8 -> int32; issued by ONE thread for the CTA
__device__ __forceinline__ void umma_i8(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
        "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}
// arrive on an mbarrier once every tcgen05.mma issued so far by this thread has completed (implies fence::before_thread_sync)
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// 8 consecutive 32-bit columns of this thread's TMEM lane (lane = 32 * (warp % 4) + lane id)
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, int32_t (&v)[8]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
                 : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
// 32 consecutive columns (four genes' limbs)
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, int32_t (&v)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, "
        "%18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
          "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]),
          "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
          "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
// 1 / d to ~1 ulp (MUFU.RCP64H seed + two Newton steps), d positive and finite; anything else gives a value the caller's
// `den > 0 && t2 < inf` test rejects or accepts correctly (0 -> inf, inf -> 0, NaN -> NaN)
__device__ __forceinline__ double um_rcp(double d) {
    double x;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(x) : "d"(d));
    double e = fma(-d, x, 1.0);
    x = fma(x, e, x);
    e = fma(-d, x, 1.0);
    return fma(x, e, x);
}
// shared-memory operand descriptor: K-major rows of 128 bytes, 128-byte swizzle, 8-row groups 1024 bytes apart
// (cute::UMMA::SmemDescriptor: start >> 4 at [0,14), LBO >> 4 at [16,30) = 1, SBO >> 4 at [32,46) = 64, version 1 at
// [46,48), layout SWIZZLE_128B = 2 at [61,64)); advancing K by 32 int8 inside the swizzle atom = + 32 bytes = + 2
__device__ __forceinline__ uint64_t umma_desc(uint32_t smem_addr) {
    return uint64_t((smem_addr >> 4) & 0x3FFFu) | (uint64_t(1) << 16) | (uint64_t(64) << 32) | (uint64_t(1) << 46) |
           (uint64_t(2) << 61);
}
// instruction descriptor (cute::UMMA::InstrDescriptor): D = S32 (2 at [4,6)), A = B = signed int8 (1 at [7,10), [10,13)),
// both K-major, N >> 3 at [17,23), M >> 4 at [24,29)
constexpr uint32_t kUmIdesc = (2u << 4) | (1u << 7) | (1u << 10) | (uint32_t(kUmTileN >> 3) << 17) | (uint32_t(kUmTileS >> 4) << 24);

// =================================================================================================
// EU1
// =================================================================================================
#ifndef STOAT_UM_EPI_WARPS
#define STOAT_UM_EPI_WARPS 8
#endif
constexpr int kUmStages = 4;
constexpr int kUmStageBytes = kUmABytes + kUmBBytes;  // 49,152
constexpr int kUmEpiWarps = STOAT_UM_EPI_WARPS;                        // two warps per TMEM lane quarter, each takes half of a tile's genes
constexpr int kUmEpiThreads = kUmEpiWarps * 32;
constexpr int kUmProducerWarp = kUmEpiWarps, kUmMmaWarp = kUmEpiWarps + 1;
constexpr int kUmThreads = (kUmEpiWarps + 2) * 32;    // warps 0-7 epilogue, 8 producer, 9 MMA

struct UmSmem {
    uint64_t full[kUmStages], empty[kUmStages], tfull[2], tempty[2];
    uint32_t tmem_base, pad_;
    EqConst k;
    EqtlRec rec[kUmTileS];
    double gconst[2][kUmTileG][3];  // per gene of the tile in flight: sum yc, sum yc^2, limb scale (one buffer per accumulator)
    // the other gene parts' best t^2 / gene / count per slot, merged by part 0's thread
    double half_key[kUmEpiWarps / 4][kUmTileS];
    uint32_t half_gene[kUmEpiWarps / 4][kUmTileS], half_cnt[kUmEpiWarps / 4][kUmTileS];
};
constexpr size_t kUmStageOff = (sizeof(UmSmem) + 1023) / 1024 * 1024;
size_t eqtl_umma_smem_bytes() { return kUmStageOff + size_t(kUmStages) * kUmStageBytes + 1024; }

__global__ void __launch_bounds__(kUmThreads, 1) eqtl_dense_umma_kernel(const EqtlDense e, const EqtlUmma u) {
    extern __shared__ uint8_t um_smem_raw[];
    // the swizzle pattern is a function of the ADDRESS: tiles must sit on 1024-byte boundaries
    uint8_t* base = um_smem_raw + ((1024u - (smem_u32(um_smem_raw) & 1023u)) & 1023u);
    UmSmem& sm = *reinterpret_cast<UmSmem*>(base);
    uint8_t* stages = base + kUmStageOff;
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // CTA order: panels of `panel` slot tiles; inside a panel the slot tile runs fastest, then the gene group.  The ~148 CTAs in
    // flight are then `panel` slot tiles x a few gene groups: their A images (panel x 655 KB at C4) and B groups stay L2 resident
    // while the panel passes over every gene group, instead of 148 different A tiles competing with one B group.
    const uint32_t n_st = uint32_t(e.s_pad / kUmTileS), panel = u.panel ? u.panel : 1u;
    const uint32_t per_panel = panel * e.n_groups;
    const uint32_t pn = blockIdx.x / per_panel, in_p = blockIdx.x % per_panel;
    const uint32_t rows_here = min(panel, n_st - pn * panel);  // the last panel may be short
    const uint32_t slot_tile = pn * panel + in_p % rows_here, grp = in_p / rows_here;
    if (grp >= e.n_groups) return;
    const uint64_t s0 = uint64_t(slot_tile) * kUmTileS;
    const uint32_t n_dense = u.probe_out ? uint32_t(kUmTileS) : *e.dense_count;
    if (s0 >= n_dense) return;  // whole CTA: nothing allocated yet
    const uint32_t g_begin = grp * e.genes_per_group;
    const uint32_t g_end = min(e.n_genes, g_begin + e.genes_per_group);
    const uint32_t n_gt = (g_end - g_begin + kUmTileG - 1) / kUmTileG;
    const uint32_t n_kb = u.n_kb;

    if (tid == 0) {
        for (int i = 0; i < kUmStages; ++i) {
            mbar_init(&sm.full[i], 1);
            mbar_init(&sm.empty[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&sm.tfull[i], 1);
            mbar_init(&sm.tempty[i], kUmEpiThreads);
        }
        mbar_fence_init();
        sm.k = EqConst{e.yc, e.tot, e.tot2, e.dcode, e.s_pad, e.n_pad, e.n_samples, e.p_threshold, e.z_threshold};
    }
    if (!u.probe_out)
        for (uint32_t i = tid; i < kUmTileS * (sizeof(EqtlRec) / 16); i += kUmThreads) {
            const bool live = s0 + i / (sizeof(EqtlRec) / 16) < n_dense;
            reinterpret_cast<uint4*>(sm.rec)[i] = live ? reinterpret_cast<const uint4*>(e.rec + s0)[i] : make_uint4(0, 0, 0, 0);
        }
    if (warp == kUmMmaWarp) tmem_alloc(&sm.tmem_base, 512);  // two accumulators of 256 columns: the whole tensor memory of the SM
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = sm.tmem_base;

    if (warp == kUmProducerWarp) {
        // ---------------- producer (the whole warp: lane 0 keeps the barriers, every lane copies a slice of the stage)
        // The 128-sample blocks are taken in an order rotated by the CTA's slot tile: the sum is exact in any order, and the
        // CTAs that share a gene group then ask the L2 for different blocks of it at any one time instead of all for the same.
        uint32_t it = 0;
        const uint32_t rot = u.rotate ? (slot_tile * 7u) % n_kb : 0u;
        const uint32_t piece = 2048;  // 24 lanes copy 2 KB each: 8 pieces of the A image, 16 of the B image
        for (uint32_t gt = 0; gt < n_gt; ++gt) {
            const uint8_t* b_src = u.b_img + (uint64_t(g_begin / kUmTileG + gt) * n_kb) * kUmBBytes;
            const uint8_t* a_src = u.a_img + (uint64_t(slot_tile) * n_kb) * kUmABytes;
            for (uint32_t kq = 0; kq < n_kb; ++kq, ++it) {
                const uint32_t kb = kq + rot < n_kb ? kq + rot : kq + rot - n_kb;
                const uint32_t st = it % kUmStages, round = it / kUmStages;
                if (round) mbar_wait(&sm.empty[st], (round - 1) & 1);
                uint8_t* dst = stages + size_t(st) * kUmStageBytes;
                if (lane == 0) mbar_arrive_expect_tx(&sm.full[st], kUmStageBytes);
                __syncwarp();
                if (u.split) {
                    if (lane >= uint32_t(kUmStageBytes) / piece) continue;
                    const uint32_t off = lane * piece;  // A image [0, 16 KB), B image [16 KB, 48 KB): pieces do not straddle
                    const uint8_t* src = off < uint32_t(kUmABytes) ? a_src + uint64_t(kb) * kUmABytes + off
                                                                   : b_src + uint64_t(kb) * kUmBBytes + (off - kUmABytes);
                    tma_bulk_g2s(dst + off, src, piece, &sm.full[st]);
                } else if (lane == 0) {
                    tma_bulk_g2s(dst, a_src + uint64_t(kb) * kUmABytes, kUmABytes, &sm.full[st]);
                    tma_bulk_g2s(dst + kUmABytes, b_src + uint64_t(kb) * kUmBBytes, kUmBBytes, &sm.full[st]);
                }
            }
        }
    } else if (warp == kUmMmaWarp) {
        // ---------------- MMA issuer (one thread)
        if (lane == 0) {
            uint32_t it = 0;
            for (uint32_t gt = 0; gt < n_gt; ++gt) {
                const uint32_t acc = gt & 1;
                if (gt >= 2) mbar_wait(&sm.tempty[acc], ((gt >> 1) - 1) & 1);  // the epilogue has drained this buffer
                tc_fence_after();
                const uint32_t d = tmem + acc * kUmTileN;
                for (uint32_t kb = 0; kb < n_kb; ++kb, ++it) {
                    const uint32_t st = it % kUmStages;
                    mbar_wait(&sm.full[st], (it / kUmStages) & 1);
                    tc_fence_after();
                    const uint32_t sa = smem_u32(stages + size_t(st) * kUmStageBytes);
                    const uint64_t da = umma_desc(sa), db = umma_desc(sa + kUmABytes);
#pragma unroll
                    for (uint32_t k = 0; k < kUmKB / 32; ++k) umma_i8(d, da + 2 * k, db + 2 * k, kUmIdesc, (kb | k) ? 1u : 0u);
                    umma_commit(&sm.empty[st]);  // the stage is free once these MMAs have read it
                }
                umma_commit(&sm.tfull[acc]);  // the gene tile's accumulator is complete
            }
        }
    } else {
        // ---------------- epilogue: thread = slot (TMEM lane), all gene tiles of the group
        // The per-pair tail of linear_regression for X = [1, d] (stats_test.cpp:455-489, as eq_pair_stat in eqtl_pair.cuh) with
        // the slot's record in registers and the gene's constants in shared memory: with four warps per SM nothing hides a
        // dependent global load per pair.  Ranking is by t^2 = b1^2 / (inv11 * rss / df) (one division, no square root); the
        // exact Student-t tail only for pairs beyond the normal-tail bound of the threshold.
        const uint32_t sl = tid & 127u;       // slot = TMEM lane; a warp reaches the lanes 32 (warp % 4) .. + 31
        const uint32_t ghalf = tid >> 7;      // which part of a tile's genes (kUmEpiWarps / 4 parts)
        constexpr uint32_t kParts = kUmEpiWarps / 4, kPartGenes = kUmTileG / kParts;
        const EqtlRec rr = sm.rec[sl];
        const bool live = rr.state == 1;
        const double n_used = double(rr.n_used);
        int df_res = int(rr.n_used) - 2 + 1;  // stats_test.cpp:467 (sic)
        if (df_res < 1) df_res = 1;
        const double dfd = double(df_res), z2 = sm.k.z_threshold * sm.k.z_threshold;
        const bool z_all = !(sm.k.z_threshold > 0.0);  // a threshold of 1 or more: every |t| has to be looked at
        const double inf = __longlong_as_double(0x7ff0000000000000LL), inv_n = 1.0 / n_used;
        double bk2 = -1.0;
        uint32_t bg = 0xFFFFFFFFu, cnt = 0;
        for (uint32_t gt = 0; gt < n_gt; ++gt) {
            const uint32_t acc = gt & 1;
            const uint32_t g_tile = g_begin + gt * kUmTileG;
            if (!u.probe_out && lane < uint32_t(kUmTileG) / kUmEpiWarps) {  // this warp stages the constants of its share of the genes
                const uint32_t gl = warp * (kUmTileG / kUmEpiWarps) + lane, g = min(g_tile + gl, e.n_genes - 1);
                sm.gconst[acc][gl][0] = e.tot[g];
                sm.gconst[acc][gl][1] = e.tot2[g];
                sm.gconst[acc][gl][2] = u.scale[g];
            }
            named_bar_sync_um(1, kUmEpiThreads);
            mbar_wait(&sm.tfull[acc], (gt >> 1) & 1);
            tc_fence_after();
            const uint32_t taddr = tmem + (((warp & 3u) * 32u) << 16) + acc * kUmTileN + ghalf * (kPartGenes * 8);
#pragma unroll 1
            for (uint32_t gb = 0; gb < kPartGenes / 4; ++gb) {
                int32_t c[32];
                tmem_ld32(taddr + gb * 32, c);  // warp-collective: every lane takes part, whatever its slot's state
                const uint32_t gl0 = ghalf * kPartGenes + gb * 4;  // first of the four genes, within the tile
                if (u.probe_out) {
#pragma unroll
                    for (int l = 0; l < 32; ++l) u.probe_out[(uint64_t(sl) * n_gt * kUmTileG + gt * kUmTileG + gl0) * 8 + l] = c[l];
                    continue;
                }
                if (!live) continue;
                double sy[4], syy[4], sdy[4], b0[4], b1[4], rss[4], k2[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    sy[j] = sm.gconst[acc][gl0 + j][0];
                    syy[j] = sm.gconst[acc][gl0 + j][1];
                }
                for (uint32_t q = 0; q < rr.n_unused; ++q) {  // samples on no path are not rows of the table
                    const uint32_t iu = rr.unused[q];
                    double y[4];
#pragma unroll
                    for (int j = 0; j < 4; ++j)  // four independent loads in flight
                        y[j] = e.yc[uint64_t(min(g_tile + gl0 + j, e.n_genes - 1)) * e.n_pad + iu];
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        sy[j] -= y[j];
                        syy[j] = fma(-y[j], y[j], syy[j]);
                    }
                }
                // the four genes in lock step, stage by stage: four independent dependency chains for the FP64 pipe
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    // sum_l C_l 128^-(l+1): every product is exact (a power of two times an integer below 2^21), smallest first
                    double v = 0.0;
#pragma unroll
                    for (int l = 7; l >= 0; --l) v = fma(double(c[8 * j + l]), __hiloint2double((1023 - 7 * (l + 1)) << 20, 0), v);
                    sdy[j] = 0.5 * sm.gconst[acc][gl0 + j][2] * v;
                }
                bool redo = false;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    b0[j] = rr.inv00 * sy[j] + rr.inv01 * sdy[j];
                    b1[j] = rr.inv01 * sy[j] + rr.inv11 * sdy[j];
                    const double bxy = b0[j] * sy[j] + b1[j] * sdy[j];
                    const double bxxb = b0[j] * (n_used * b0[j] + rr.sd * b1[j]) + b1[j] * (rr.sd * b0[j] + rr.sdd * b1[j]);
                    rss[j] = syy[j] - 2.0 * bxy + bxxb;
                    const double tss = syy[j] - sy[j] * sy[j] * inv_n;  // only decides which form of rss is taken
                    if (!(rss[j] > 1e-6 * tss)) {
                        rss[j] = -1.0;  // marker: the moment expansion has lost its digits (near-perfect fit)
                        redo = true;
                    }
                }
                if (redo) {
#pragma unroll 1
                    for (int j = 0; j < 4; ++j)
                        if (rss[j] == -1.0 && g_tile + gl0 + j < g_end)
                            rss[j] = eq_rss_explicit(e.yc + uint64_t(g_tile + gl0 + j) * e.n_pad, e.dcode + (s0 + sl) * 16, e.s_pad,
                                                     e.n_samples, &sm.rec[sl], b0[j], b1[j]);
                }
                bool look = z_all;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const double den = rr.inv11 * (rss[j] / dfd);
                    const double t2 = (b1[j] * b1[j]) * um_rcp(den);  // ~1 ulp; |t| is reported to 1e-9
                    const bool valid = g_tile + gl0 + j < g_end;
                    const bool fin = den > 0.0 && t2 < inf;  // t finite (:480-482), else p = 1
                    k2[j] = !valid ? -2.0 : fin ? t2 : -0.5;  // -0.5: invalid t, ranks as t = 0
                    look = look || k2[j] > z2 || (k2[j] == -0.5 && 1.0 < sm.k.p_threshold);
                }
                if (look) {  // rare at a genome-wide threshold: the exact Student-t tail decides
#pragma unroll 1
                    for (int j = 0; j < 4; ++j) {
                        if (k2[j] == -2.0) continue;
                        bool below;
                        if (k2[j] >= 0.0) below = (z_all || k2[j] > z2) && 2.0 * student_t_upper_tail(dfd, sqrt(k2[j])) < sm.k.p_threshold;
                        else below = 1.0 < sm.k.p_threshold;
                        cnt += below ? 1u : 0u;
                    }
                }
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const double kk = k2[j] == -0.5 ? 0.0 : k2[j];  // (-2: not a gene, never the best)
                    const uint32_t g = g_tile + gl0 + j;
                    if (kk > bk2 || (kk == bk2 && kk >= 0.0 && g < bg)) {
                        bk2 = kk;
                        bg = g;
                    }
                }
            }
            tc_fence_before();
            mbar_arrive(&sm.tempty[acc]);
        }
        if (!u.probe_out) {
            // the two gene halves of a slot: the upper one hands over through shared memory
            if (ghalf != 0) {
                sm.half_key[ghalf][sl] = bk2;
                sm.half_gene[ghalf][sl] = bg;
                sm.half_cnt[ghalf][sl] = cnt;
            }
            named_bar_sync_um(1, kUmEpiThreads);
            if (ghalf == 0) {
                for (uint32_t h = 1; h < kParts; ++h) {
                    const double ok = sm.half_key[h][sl];
                    const uint32_t og = sm.half_gene[h][sl];
                    cnt += sm.half_cnt[h][sl];
                    if (ok > bk2 || (ok == bk2 && og < bg)) {
                        bk2 = ok;
                        bg = og;
                    }
                }
                const uint64_t o = uint64_t(grp) * e.s_pad + s0 + sl;
                e.part_key[o] = bk2 < 0.0 ? -1.0 : sqrt(bk2);  // |t| of the best gene, as the FP64 kernel reports it
                e.part_gene[o] = bg;
                e.part_cnt[o] = cnt;
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == kUmMmaWarp) tmem_dealloc(tmem, 512);
}

cudaError_t launch_eqtl_umma_prepare_expression(const double* yc, uint32_t n_genes, uint32_t n_pad, double* scale,
                                                uint8_t* b_img, cudaStream_t st) {
    if (n_genes == 0) return cudaSuccess;
    const uint32_t n_kb = n_pad / kUmKB;
    eqtl_gene_scale_kernel<<<n_genes, 256, 0, st>>>(yc, n_pad, scale);
    cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess) return err;
    const uint64_t g_pad = (uint64_t(n_genes) + kUmTileG - 1) / kUmTileG * kUmTileG;
    const uint64_t threads = g_pad * (n_pad / 16);
    eqtl_limbs_kernel<<<unsigned((threads + 255) / 256), 256, 0, st>>>(yc, scale, n_genes, n_pad, n_kb, b_img);
    return cudaGetLastError();
}

cudaError_t launch_eqtl_dense_umma(const EqtlDense& e, const EqtlUmma& u, cudaStream_t st) {
    if (e.s_pad == 0 || e.n_genes == 0) return cudaSuccess;
    cudaError_t err;
    if (!u.probe_out) {
        const uint64_t threads = e.s_pad * uint64_t(e.n_kst) * 4;
        eqtl_expand_codes_kernel<<<unsigned((threads + 255) / 256), 256, 0, st>>>(e.dcode, e.s_pad, e.n_kst, u.n_kb, e.dense_count,
                                                                                 u.a_img);
        if ((err = cudaGetLastError()) != cudaSuccess) return err;
    }
    const int smem = int(eqtl_umma_smem_bytes());
    err = cudaFuncSetAttribute(eqtl_dense_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (err != cudaSuccess) return err;
    const uint64_t n_st = e.s_pad / kUmTileS, panel = u.panel ? u.panel : 1;
    const uint64_t n_panels = (n_st + panel - 1) / panel;
    eqtl_dense_umma_kernel<<<unsigned(n_panels * panel * e.n_groups), kUmThreads, smem, st>>>(e, u);
    return cudaGetLastError();
}

}  // namespace stoat
