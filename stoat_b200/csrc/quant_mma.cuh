// quant_mma.cuh — K3M, the X^T y / class-count pass of the quantitative test for two-path snarls as an EXACT integer
// contraction (included by kernels_quant.cu, namespace stoat).
//
// Why: K3 (one warp per snarl, phenotype in shared memory) and K3R (one CTA per snarl, phenotype in registers) both need
// several FP64-pipe instructions per (snarl, sample) and stay issue / latency bound at ~0.15 of the HBM roofline.  The
// dosage of a sample on path 0 takes the values 0, 1/2, 1, i.e. 2d is a 2-bit integer, and the centred phenotype can be
// split ONCE (host, set_quantitative_phenotype) into 8 signed 7-bit slices of a common fixed-point scale,
//       y_i = scale * sum_t q[i][t] * 128^-(t+1),   |residual| <= 2^-57 scale          (an Ozaki split)
// so  sum_i 2d_i y_i = scale * sum_t 128^-(t+1) * C_t  with  C_t = sum_i (2 d_i) q[i][t]  an exact INT32 dot product.
// C[snarl][t] is a 16 x 8 x 32 integer MMA (mma.sync.m16n8k32.s8, SASS IMMA): 16 snarls share the 8 slices of 32 samples,
// the sample -> k permutation is free, and the dosage bytes come straight from the class bit masks with one shift and one
// AND per FOUR samples.  Per (snarl, sample) the kernel issues ~0.07 warp instructions (K3R: 0.33, K3: 0.64); the result
// is more accurate than an FP64 summation (no rounding inside the dot products) and bit-reproducible.
//
// Work split: a CTA (8 warps) takes a tile of 16 consecutive snarls; their distinct rows are TMA-staged into 16 slots of
// a stage (two stages, ~80 KB each at H = 10k: one in flight while the other is consumed).  Warp w takes the 16-byte
// chunks w, w + 8, ... of the rows; lane (g = lane / 4, t = lane % 4) takes word t of a chunk for the snarls g and g + 8
// of the tile: path words (2 x 2 LDS + AND), class masks (dosage_masks.cuh), packed popcounts, 4 dosage registers, two
// MMAs against the slice bytes of the chunk (LDS.64 from the CTA's copy of the slice table).  After its chunks a warp
// adds its 16 x 8 accumulators and packed counts into the tile's shared-memory sums with integer atomics (order free,
// hence deterministic); the last warp to arrive turns them into the slice partial sums K4 consumes (quant_reg.cuh:
// LeanPartials) — the same hand-over a sample-sharded run all-reduces.
//
// Scope: np == 2 and both paths with 1 or 2 checked edges, no sample on three path-haplotypes, H < 65536.  Everything
// else goes to the slow list (K3's warp-per-snarl kernel).  Samples on neither path are rare: their phenotype sums are
// accumulated in FP64 from global memory by the lanes that meet them.
#pragma once

constexpr int kMmaTile = 16;    // snarls per tile (the M of the MMA)
constexpr int kMmaWarps = 8;
constexpr int kMmaSlices = 8;   // phenotype slices (the N of the MMA)

struct K3MParams {
    K3Params k;
    const uint8_t* yslices;   // [chunk][2][32 lanes][8 B]: B fragments of the two MMAs of a chunk
    const double* yc;         // centred phenotype, sample order (rare path: samples on neither path)
    double y_scale;           // y = y_scale * sum_t q_t 128^-(t+1)
    uint32_t chunk_first, chunk_count;  // per slice (blockIdx.y): chunks [y * chunk_count, ...) of pitch / 16
    uint32_t row_stride;      // bytes between the rows of a slot (= bytes of a row slice)
    uint32_t slot_bytes;      // bytes between the slots of a stage (== 16 mod 128: the 8 groups hit different banks)
    uint32_t* slow_list;
    uint32_t* slow_count;
    LeanPartials part;
};

__device__ __forceinline__ void mma_s8_16832(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                             uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// per-tile sums in shared memory
struct MmaTileSums {
    int acc[kMmaTile][kMmaSlices];       // C_t
    uint32_t cnt[kMmaTile][3];           // al0 | al1 << 16, n_used | n_full0 << 16, n_half
    uint32_t flags[kMmaTile];            // bit 0: a sample on three path-haplotypes (-> slow list)
    double un[kMmaWarps][kMmaTile][2];   // per warp: sum y, sum y^2 over the samples on neither path
    uint32_t arrived;
    uint32_t pad[3];
};

// what a lane knows about one of its two snarls
struct MmaSnarl {
    uint32_t off[4];  // byte offsets of the rows path 0 ANDs (off[0], off[1]) and path 1 ANDs (off[2], off[3]) in the stage
    uint32_t live;    // 0xFFFFFFFF when the snarl is done here, 0 when it goes to the slow list / is past the end
};

// stage a tile: slot k of the stage <- path program + distinct rows of snarl tile * 16 + k (warp-collective, warp 0)
__device__ __forceinline__ void mma_issue_tile(const K3MParams& q, uint64_t tile, uint8_t* stage, uint64_t* bar,
                                               uint32_t slice_off, uint32_t slice_bytes, uint32_t lane, uint64_t policy) {
    const K3Params& p = q.k;
    const uint64_t s0 = tile * kMmaTile;
    // record words of the 16 snarls: lane = 8 * (k % 4) + word, pass = k / 4
    uint32_t rw[4];
#pragma unroll
    for (int pass = 0; pass < 4; ++pass) {
        const uint64_t s = s0 + uint64_t(pass) * 4 + (lane >> 3);
        rw[pass] = (s < p.n_snarls) ? __ldg(reinterpret_cast<const uint32_t*>(p.rec + s) + (lane & 7)) : 0u;
    }
#pragma unroll
    for (int k = 0; k < kMmaTile; ++k) {
        uint32_t w = __shfl_sync(0xffffffffu, rw[k >> 2], (k & 3) * 8 + (lane & 7));
        // snarls this kernel does not take (not two paths / more than 4 distinct rows): program only, no rows
        if (lane == 7 && ((w >> 16) != 2u || (w & 0xFFu) > 4u)) w &= ~0xFFu;
        if (s0 + k < p.n_snarls) {
            issue_stage_slice(w, lane, stage + size_t(k) * q.slot_bytes, bar, p.rows, p.blob, p.pitch, slice_off,
                              slice_bytes, q.row_stride, policy);
        } else if (lane == 0) {
            mbar_arrive(bar);  // no such snarl: keep the arrival count of the phase at 16
        }
    }
}

// =================================================================================================
// K3M kernel.  grid = (CTAs per slice, slices); CTA b of a slice takes tiles b, b + gridDim.x, ...
// dynamic shared memory: [6 mbarriers | pad 128][2 x MmaTileSums][slice table of this slice][2 stages x 16 slots]
// =================================================================================================
__global__ void __launch_bounds__(kMmaWarps * 32, 1) dosage_moments_mma_kernel(const K3MParams q) {
    extern __shared__ __align__(128) uint8_t smem[];
    const K3Params& p = q.k;
    const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
    uint64_t* full = reinterpret_cast<uint64_t*>(smem);
    uint64_t* empty = full + 2;
    uint64_t* bfree = empty + 2;
    MmaTileSums* sums = reinterpret_cast<MmaTileSums*>(smem + 128);
    const uint32_t sums_bytes = (2u * uint32_t(sizeof(MmaTileSums)) + 127u) & ~127u;
    uint8_t* ysl = smem + 128 + sums_bytes;
    const uint32_t slice = blockIdx.y;
    const uint32_t c_first = slice * q.chunk_count;
    const uint32_t c_len = min(q.chunk_count, p.pitch / 16 - c_first);
    uint8_t* stages = ysl + ((q.chunk_count * 512u + 127u) & ~127u);
    const uint32_t stage_bytes = kMmaTile * q.slot_bytes;
    const uint32_t slice_off = c_first * 16, slice_bytes = c_len * 16;

    if (tid == 0) {
        for (int i = 0; i < 2; ++i) {
            mbar_init(&full[i], kMmaTile);
            mbar_init(&empty[i], kMmaWarps);
            mbar_init(&bfree[i], 1);
        }
        mbar_fence_init();
    }
    for (uint32_t i = tid; i < 2u * sizeof(MmaTileSums) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(sums)[i] = 0u;
    {
        const uint4* src = reinterpret_cast<const uint4*>(q.yslices + size_t(c_first) * 512);
        uint4* dst = reinterpret_cast<uint4*>(ysl);
        for (uint32_t i = tid; i < c_len * 32u; i += blockDim.x) dst[i] = __ldg(src + i);
    }
    __syncthreads();

    const uint64_t n_tiles = (p.n_snarls + kMmaTile - 1) / kMmaTile;
    const uint64_t stride = gridDim.x;
    const uint64_t policy = make_evict_first_policy();
    if (warp == 0) {
        uint64_t tl = blockIdx.x;
        for (uint32_t st = 0; st < 2 && tl < n_tiles; ++st, tl += stride)
            mma_issue_tile(q, tl, stages + size_t(st) * stage_bytes, &full[st], slice_off, slice_bytes, lane, policy);
    }

    uint32_t it = 0;
    for (uint64_t tile = blockIdx.x; tile < n_tiles; tile += stride, ++it) {
        const uint32_t st = it & 1u, par = (it >> 1) & 1u;
        // ---- warp 0: once everybody has left the previous tile's stage, refill it with the tile after this one ----
        if (warp == 0 && it >= 1u && tile + stride < n_tiles) {
            mbar_wait(&empty[st ^ 1u], ((it - 1u) >> 1) & 1u);
            mma_issue_tile(q, tile + stride, stages + size_t(st ^ 1u) * stage_bytes, &full[st ^ 1u], slice_off, slice_bytes,
                           lane, policy);
        }
        mbar_wait(&full[st], par);
        const uint8_t* stage = stages + size_t(st) * stage_bytes;
        // ---- this lane's two snarls: slots g and g + 8 ----------------------------------------------------------
        MmaSnarl sn[2];
        uint32_t info[2];  // bit 1: done here, bit 2: has paths (np != 0)
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const uint32_t slot = g + 8u * h;
            const uint32_t base = slot * q.slot_bytes;
            const uint4 hdr = *reinterpret_cast<const uint4*>(stage + base);  // np, class | p0 | codes, locals ...
            const uint32_t np = hdr.x & 0xFFu, code0 = hdr.z & 0xFFu, code1 = (hdr.z >> 8) & 0xFFu;
            const bool in_range = tile * kMmaTile + slot < p.n_snarls;
            const bool ok = in_range && np == 2u && (code0 == 1u || code0 == 2u) && (code1 == 1u || code1 == 2u);
            const uint64_t loc = (uint64_t(hdr.w) << 16) | (hdr.z >> 16);  // bytes 10.. of the program
            const uint32_t l0 = uint32_t(loc) & 0xFFu;
            const uint32_t l1 = (code0 == 2u) ? (uint32_t(loc >> 8) & 0xFFu) : l0;
            const uint64_t loc1 = loc >> (8u * (code0 & 3u));
            const uint32_t l2 = uint32_t(loc1) & 0xFFu;
            const uint32_t l3 = (code1 == 2u) ? (uint32_t(loc1 >> 8) & 0xFFu) : l2;
            const uint32_t rows0 = base + kStageBlobBytes;
            sn[h].off[0] = rows0 + (ok ? l0 : 0u) * q.row_stride;
            sn[h].off[1] = rows0 + (ok ? l1 : 0u) * q.row_stride;
            sn[h].off[2] = rows0 + (ok ? l2 : 0u) * q.row_stride;
            sn[h].off[3] = rows0 + (ok ? l3 : 0u) * q.row_stride;
            sn[h].live = ok ? 0xFFFFFFFFu : 0u;
            info[h] = (ok ? 2u : 0u) | ((in_range && np != 0u) ? 4u : 0u);
        }
        int c[4] = {0, 0, 0, 0};
        uint32_t cA[2] = {0, 0}, cB[2] = {0, 0}, cC[2] = {0, 0}, nd[2] = {0, 0};
        double su[2] = {0.0, 0.0}, suu[2] = {0.0, 0.0};
        bool saw_unused = false;
        for (uint32_t ch = warp; ch < c_len; ch += kMmaWarps) {
            const uint32_t woff = ch * 16 + t * 4;
            const uint32_t word = (c_first + ch) * 4 + t;
            const uint32_t valid = valid_word(word, p.n_hap) & 0x55555555u;
            const uint2 bm0 = *reinterpret_cast<const uint2*>(ysl + (size_t(ch * 2 + 0) * 32 + lane) * 8);
            const uint2 bm1 = *reinterpret_cast<const uint2*>(ysl + (size_t(ch * 2 + 1) * 32 + lane) * 8);
            uint32_t r[2][4];
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const uint32_t a = *reinterpret_cast<const uint32_t*>(stage + sn[h].off[0] + woff) &
                                   *reinterpret_cast<const uint32_t*>(stage + sn[h].off[1] + woff) & sn[h].live;
                const uint32_t b = *reinterpret_cast<const uint32_t*>(stage + sn[h].off[2] + woff) &
                                   *reinterpret_cast<const uint32_t*>(stage + sn[h].off[3] + woff) & sn[h].live;
                const LeanMasks m = lean_masks(a, b);
                cA[h] += __popc(a) | (__popc(b) << 16);
                cB[h] += __popc(m.used) | (__popc(m.full0) << 16);
                cC[h] += __popc(m.half);
                nd[h] |= m.third | m.tthird;
                uint32_t un = ~m.used & valid & sn[h].live;
                if (un) {  // rare: samples on neither path — their phenotype comes from global memory (L2)
                    saw_unused = true;
                    while (un) {
                        const int bit = __ffs(un) - 1;
                        un &= un - 1;
                        const double yv = __ldg(q.yc + size_t(word) * 16 + (bit >> 1));
                        su[h] += yv;
                        suu[h] = fma(yv, yv, suu[h]);
                    }
                }
                const uint32_t fld = m.half | (m.full0 << 1);  // 2-bit field per sample: 2 * dosage
                r[h][0] = fld & 0x03030303u;          // samples 0, 4, 8, 12 of the word, one per byte
                r[h][1] = (fld >> 2) & 0x03030303u;   // samples 1, 5, 9, 13
                r[h][2] = (fld >> 4) & 0x03030303u;   // samples 2, 6, 10, 14
                r[h][3] = (fld >> 6) & 0x03030303u;   // samples 3, 7, 11, 15
            }
            mma_s8_16832(c, r[0][0], r[1][0], r[0][1], r[1][1], bm0.x, bm0.y);
            mma_s8_16832(c, r[0][2], r[1][2], r[0][3], r[1][3], bm1.x, bm1.y);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[st]);  // this warp is done with the stage

        // ---- add this warp's sums into the tile's sums (integer atomics: order free) -----------------------------
        MmaTileSums& S = sums[st];
        if (it >= 2u) mbar_wait(&bfree[st], par ^ 1u);  // the tile two back (same buffer) has been finished
#pragma unroll
        for (int h = 0; h < 2; ++h) {  // the 4 lanes of a group hold the words of the same two snarls
            cA[h] += __shfl_xor_sync(0xffffffffu, cA[h], 1);
            cB[h] += __shfl_xor_sync(0xffffffffu, cB[h], 1);
            cC[h] += __shfl_xor_sync(0xffffffffu, cC[h], 1);
            nd[h] |= __shfl_xor_sync(0xffffffffu, nd[h], 1);
            cA[h] += __shfl_xor_sync(0xffffffffu, cA[h], 2);
            cB[h] += __shfl_xor_sync(0xffffffffu, cB[h], 2);
            cC[h] += __shfl_xor_sync(0xffffffffu, cC[h], 2);
            nd[h] |= __shfl_xor_sync(0xffffffffu, nd[h], 2);
        }
        if (__any_sync(0xffffffffu, saw_unused)) {
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                su[h] += __shfl_xor_sync(0xffffffffu, su[h], 1);
                suu[h] += __shfl_xor_sync(0xffffffffu, suu[h], 1);
                su[h] += __shfl_xor_sync(0xffffffffu, su[h], 2);
                suu[h] += __shfl_xor_sync(0xffffffffu, suu[h], 2);
            }
        }
        atomicAdd(&S.acc[g][2 * t], c[0]);
        atomicAdd(&S.acc[g][2 * t + 1], c[1]);
        atomicAdd(&S.acc[g + 8][2 * t], c[2]);
        atomicAdd(&S.acc[g + 8][2 * t + 1], c[3]);
        if (t == 0) {
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const uint32_t slot = g + 8u * h;
                atomicAdd(&S.cnt[slot][0], cA[h]);
                atomicAdd(&S.cnt[slot][1], cB[h]);
                atomicAdd(&S.cnt[slot][2], cC[h]);
                atomicOr(&S.flags[slot], info[h] | (nd[h] ? 1u : 0u));
                S.un[warp][slot][0] = su[h];
                S.un[warp][slot][1] = suu[h];
            }
        }
        __syncwarp();
        uint32_t last = 0;
        if (lane == 0) {
            __threadfence_block();
            last = (atomicAdd(&S.arrived, 1u) == uint32_t(kMmaWarps) - 1u) ? 1u : 0u;
        }
        last = __shfl_sync(0xffffffffu, last, 0);
        if (last) {  // ---- the last warp to arrive finishes the tile: lane <-> slot ------------------------------
            __threadfence_block();
            if (lane < uint32_t(kMmaTile)) {
                const uint64_t s = tile * kMmaTile + lane;
                const uint32_t fl = S.flags[lane];
                if (s < p.n_snarls) {
                    if (fl & 2u) {
                        // Horner over the slices, smallest first: sum_t 128^-(t+1) C_t
                        double hsum = 0.0;
#pragma unroll
                        for (int k = kMmaSlices - 1; k >= 0; --k) hsum = (hsum + double(S.acc[lane][k])) * (1.0 / 128.0);
                        const double xty1 = 0.5 * q.y_scale * hsum;  // d = (2d) / 2
                        double su_un = 0.0, suu_un = 0.0;
#pragma unroll
                        for (int w = 0; w < kMmaWarps; ++w) {
                            su_un += S.un[w][lane][0];
                            suu_un += S.un[w][lane][1];
                        }
                        const uint32_t pa = S.cnt[lane][0], pb = S.cnt[lane][1], pc = S.cnt[lane][2];
                        uint32_t* pi = q.part.i + size_t(slice) * 7 * q.part.stride + s;
                        double* pd = q.part.d + size_t(slice) * 3 * q.part.stride + s;
                        pi[0] = pa & 0xFFFFu;
                        pi[q.part.stride] = pa >> 16;
                        pi[2 * q.part.stride] = pb & 0xFFFFu;
                        pi[3 * q.part.stride] = pb >> 16;
                        pi[4 * q.part.stride] = pc;
                        pi[5 * q.part.stride] = 0u;
                        pi[6 * q.part.stride] = (fl & 1u) ? kLeanReplayMark : 0u;  // K4 sends the snarl to the general path
                        pd[0] = xty1;
                        pd[q.part.stride] = su_un;
                        pd[2 * q.part.stride] = suu_un;
                        if (slice == 0) p.mom[s].state = 2;  // K4 builds the moments from the partial sums
                    } else if (slice == 0) {
                        p.mom[s].state = 0;
                        if (fl & 4u) {  // np == 0: prep already put the snarl on the general list
                            const uint32_t slotx = atomicAdd(q.slow_count, 1u);
                            q.slow_list[slotx] = uint32_t(s);
                        }
                    }
                }
            }
            __syncwarp();
            // clear the integer sums for the tile that uses this buffer next
            for (uint32_t i = lane; i < kMmaTile * kMmaSlices; i += 32) (&S.acc[0][0])[i] = 0;
            for (uint32_t i = lane; i < kMmaTile * 3; i += 32) (&S.cnt[0][0])[i] = 0u;
            if (lane < uint32_t(kMmaTile)) S.flags[lane] = 0u;
            __syncwarp();
            if (lane == 0) {
                S.arrived = 0u;
                mbar_arrive(&bfree[st]);  // (release: the clears are visible before anybody passes the wait above)
            }
        }
    }
}
