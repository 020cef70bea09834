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
constexpr int kMmaWarps = 8;    // warps per CTA (they share the slice table, nothing else)
constexpr int kMmaSlices = 8;   // phenotype slices (the N of the MMA)
constexpr int kMmaMaxRounds = 24;  // rounds (4 chunks = 64 B of every row) per haplotype slice: 48 KB of slice table

// per snarl: the four matrix rows path 0 (r[0] & r[1]) and path 1 (r[2] & r[3]) AND; r[0] == kLeanNo: not done here
// (r[1] != 0: the snarl has paths and goes to the slow list)
constexpr uint32_t kLeanNo = 0xFFFFFFFFu;
constexpr uint32_t kLeanAll = 0xFFFFFFFEu;  // rec2x: a row that every haplotype has (a path without a checked edge)

struct K3MParams {
    K3Params k;
    const uint4* lean_rec;    // [n_snarls]
    const uint8_t* yslices;   // [round][word 4][mma 2][32 lanes][8 B]: B fragments (see abi.cu, set_quantitative_phenotype)
    const double* yc;         // centred phenotype, sample order (rare path: samples on neither path)
    double y_scale;           // y = y_scale * sum_t q_t 128^-(t+1)
    uint32_t rounds_per_slice;  // blockIdx.y covers rounds [y * rounds_per_slice, ...)
    uint32_t n_rounds;          // ceil(chunks / 4)
    uint32_t prefetch;          // 1: L2 prefetch of the rows four to eight rounds ahead
    uint32_t* slow_list;
    uint32_t* slow_count;
    LeanPartials part;
    uint4* rec2x;        // nullptr: snarls with non-dyadic dosages are marked for the general path instead
    uint32_t* count2x;      // records lean_rec_kernel listed (fixed for a staging)
    uint32_t* count2x_dyn;  // records K3M appended behind them during this step
    uint32_t table_in_global;  // K3M3 / K3M2x: the slice table does not fit shared memory (> 24 rounds = 12,288 haplotypes),
                               // the B fragments are read from global memory (L1 / L2 resident: 2 KB per round)
};

// rec / path program -> lean record (one thread per snarl; 48 B in, 16 B out)
// Three-path snarls (rec3 != nullptr): every path with 1 or 2 checked edges -> a record of six rows in rec3 (compacted,
// order free: the results are stored by snarl) for dosage_moments_mma3_kernel, and nothing for the slow list.
__global__ void __launch_bounds__(256) lean_rec_kernel(const SnarlRec* __restrict__ rec, const uint8_t* __restrict__ blob,
                                                       uint64_t n_snarls, uint4* __restrict__ out, uint4* __restrict__ rec3,
                                                       uint32_t* __restrict__ count3, uint4* __restrict__ rec2x,
                                                       uint32_t* __restrict__ count2x) {
    const uint64_t s = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    const bool in = s < n_snarls;
    SnarlRec r{};
    uint4 hdr = make_uint4(0u, 0u, 0u, 0u);
    if (in) {
        r = rec[s];
        hdr = *reinterpret_cast<const uint4*>(blob + size_t(r.blob16) * 16);  // np, class | p0 | codes, locals ...
    }
    const uint32_t np = hdr.x & 0xFFu, code0 = hdr.z & 0xFFu, code1 = (hdr.z >> 8) & 0xFFu;
    const uint32_t code2 = (hdr.z >> 16) & 0xFFu;
    const bool ok3 = in && rec3 != nullptr && r.n_rows != 0 && np == 3u && (code0 == 1u || code0 == 2u) &&
                     (code1 == 1u || code1 == 2u) && (code2 == 1u || code2 == 2u);
    {
        const uint32_t votes = __ballot_sync(0xffffffffu, ok3);
        uint32_t base = 0;
        const uint32_t lane = threadIdx.x & 31u;
        if (votes != 0u) {
            if (lane == 0) base = atomicAdd(count3, uint32_t(__popc(votes)));
            base = __shfl_sync(0xffffffffu, base, 0);
        }
        if (ok3) {
            // program bytes: 8.. codes (3), 11.. locals of path 0, 1, 2
            const uint8_t* loc = blob + size_t(r.blob16) * 16 + kBlobHeaderQ + 3;
            uint32_t rows6[6];
            uint32_t off = 0;
            const uint32_t codes[3] = {code0, code1, code2};
#pragma unroll
            for (int j = 0; j < 3; ++j) {
                const uint32_t la = loc[off], lb = (codes[j] == 2u) ? loc[off + 1] : la;
                rows6[2 * j] = r.rows[la];
                rows6[2 * j + 1] = r.rows[lb];
                off += codes[j];
            }
            const uint32_t slot = base + uint32_t(__popc(votes & ((1u << lane) - 1u)));
            rec3[2 * size_t(slot)] = make_uint4(rows6[0], rows6[1], rows6[2], rows6[3]);
            rec3[2 * size_t(slot) + 1] = make_uint4(rows6[4], rows6[5], hdr.y, uint32_t(s));
        }
    }
    // Two paths of which one has no checked edge (it matches every haplotype, snarl_analyzer.cpp:343-354): every sample
    // that takes the other path sits on three or four path-haplotypes -> dosages 1/3, 1/2: the 2x kernel (rec2x != nullptr)
    const bool c0ok = code0 <= 2u, c1ok = code1 <= 2u;
    const bool ok2x = in && rec2x != nullptr && r.n_rows != 0 && np == 2u && c0ok && c1ok && (code0 == 0u) != (code1 == 0u);
    {
        const uint32_t votes = __ballot_sync(0xffffffffu, ok2x);
        uint32_t base = 0;
        const uint32_t lane = threadIdx.x & 31u;
        if (votes != 0u) {
            if (lane == 0) base = atomicAdd(count2x, uint32_t(__popc(votes)));
            base = __shfl_sync(0xffffffffu, base, 0);
        }
        if (ok2x) {
            const uint8_t* loc = blob + size_t(r.blob16) * 16 + kBlobHeaderQ + 2;
            uint32_t rows4[4];
            uint32_t off = 0;
            const uint32_t codes[2] = {code0, code1};
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                if (codes[j] == 0u) {
                    rows4[2 * j] = rows4[2 * j + 1] = kLeanAll;
                } else {
                    const uint32_t la = loc[off], lb = (codes[j] == 2u) ? loc[off + 1] : la;
                    rows4[2 * j] = r.rows[la];
                    rows4[2 * j + 1] = r.rows[lb];
                }
                off += codes[j];
            }
            const uint32_t slot = base + uint32_t(__popc(votes & ((1u << lane) - 1u)));
            rec2x[2 * size_t(slot)] = make_uint4(rows4[0], rows4[1], rows4[2], rows4[3]);
            rec2x[2 * size_t(slot) + 1] = make_uint4(0u, 0u, hdr.y, uint32_t(s));
        }
    }
    if (!in) return;
    const bool ok = r.n_rows != 0 && np == 2u && (code0 == 1u || code0 == 2u) && (code1 == 1u || code1 == 2u);
    uint4 o = make_uint4(kLeanNo, (np != 0u && !ok3 && !ok2x) ? 1u : 0u, hdr.y, 0u);
    if (ok) {
        const uint64_t loc = (uint64_t(hdr.w) << 16) | (hdr.z >> 16);  // bytes 10.. of the program: locals
        const uint32_t l0 = uint32_t(loc) & 0xFFu;
        const uint32_t l1 = (code0 == 2u) ? (uint32_t(loc >> 8) & 0xFFu) : l0;
        const uint64_t loc1 = loc >> (8u * code0);
        const uint32_t l2 = uint32_t(loc1) & 0xFFu;
        const uint32_t l3 = (code1 == 2u) ? (uint32_t(loc1 >> 8) & 0xFFu) : l2;
        o = make_uint4(r.rows[l0], r.rows[l1], r.rows[l2], r.rows[l3]);
    }
    out[s] = o;
}

__device__ __forceinline__ void mma_s8_16832(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                             uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// running sums of a lane over one tile: its two snarls (h = 0: slot g, h = 1: slot g + 8)
struct MmaLaneSums {
    int c[4];
    uint32_t cA[2], cB[2], cC[2], nd[2];  // al0 | al1 << 16, n_used | n_full0 << 16, n_half, third-class bits seen
    double su[2], suu[2];                 // sum y, sum y^2 over the samples on neither path
    uint32_t live[2];                     // 0xFFFFFFFF: the snarl is done here
    bool saw_unused;
};

// one 32-haplotype word of both snarls: class counts, dosage bytes, two MMAs
__device__ __forceinline__ void mma_word_step(const K3MParams& q, MmaLaneSums& z, const uint32_t (&a)[2],
                                              const uint32_t (&b)[2], uint32_t word, const uint8_t* bfrag) {
    const K3Params& p = q.k;
    const uint2 bm0 = *reinterpret_cast<const uint2*>(bfrag);
    const uint2 bm1 = *reinterpret_cast<const uint2*>(bfrag + 256);
    uint32_t r[2][4], used[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const LeanMasks m = lean_masks(a[h], b[h]);
        used[h] = m.used;
        z.cA[h] += __popc(a[h]) | (__popc(b[h]) << 16);
        z.cB[h] += __popc(m.used) | (__popc(m.full0) << 16);
        z.cC[h] += __popc(m.half);
        z.nd[h] |= m.third | m.tthird;
        const uint32_t fld = m.half | (m.full0 << 1);  // 2-bit field per sample: 2 * dosage
        r[h][0] = fld & 0x03030303u;          // samples 0, 4, 8, 12 of the word, one per byte
        r[h][1] = (fld >> 2) & 0x03030303u;   // samples 1, 5, 9, 13
        r[h][2] = (fld >> 4) & 0x03030303u;   // samples 2, 6, 10, 14
        r[h][3] = (fld >> 6) & 0x03030303u;   // samples 3, 7, 11, 15
    }
    mma_s8_16832(z.c, r[0][0], r[1][0], r[0][1], r[1][1], bm0.x, bm0.y);
    mma_s8_16832(z.c, r[0][2], r[1][2], r[0][3], r[1][3], bm1.x, bm1.y);
    // rare: samples on neither path of a snarl that is done here (idle chunks lie beyond n_hap: valid == 0)
    const uint32_t valid = valid_word(word, p.n_hap) & 0x55555555u;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        uint32_t un = ~used[h] & valid & z.live[h];
        if (un != 0u) {
            z.saw_unused = true;
            while (un) {
                const int bit = __ffs(un) - 1;
                un &= un - 1;
                const double yv = __ldg(q.yc + size_t(word) * 16 + (bit >> 1));
                z.su[h] += yv;
                z.suu[h] = fma(yv, yv, z.suu[h]);
            }
        }
    }
}

// rare, and kept out of line so that it does not take part in the register allocation of the K3M loop
__device__ __noinline__ void mma_push_thirds(uint4* rec2x, const uint32_t* count2x, uint32_t* count2x_dyn, QuantMoments* mom,
                                             uint32_t s, uint4 rows) {
    mom[s].state = 0;
    const uint32_t slot = *count2x + atomicAdd(count2x_dyn, 1u);
    rec2x[2 * size_t(slot)] = rows;
    rec2x[2 * size_t(slot) + 1] = make_uint4(0u, 0u, 0u, s);
}

// the 64 bytes (4 words) of chunk 4 * round + t of the four rows of both snarls
struct MmaRound {
    uint4 v[2][4];
};
__device__ __forceinline__ void mma_load_round(const K3MParams& q, const uint4 (&lr)[2], const uint32_t (&live)[2],
                                               uint32_t chunk, MmaRound& d) {
    const K3Params& p = q.k;
    const bool in = chunk < p.pitch / 16;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const uint32_t rr[4] = {lr[h].x, lr[h].y, lr[h].z, lr[h].w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            d.v[h][k] = make_uint4(0u, 0u, 0u, 0u);
            if (in && live[h])
                d.v[h][k] = ldg_stream(reinterpret_cast<const uint4*>(p.rows + size_t(rr[k]) * p.pitch + size_t(chunk) * 16));
        }
    }
}
// L2 prefetch of 256 bytes (4 rounds) of the rows of this lane group's two snarls: lane t takes row t of each.  The
// kernel's own loads are 64-byte pieces of 64 different rows per warp instruction — poor DRAM page locality when they
// miss; the prefetch turns the DRAM side into 256-byte bursts and the LDGs into L2 hits.
__device__ __forceinline__ void mma_prefetch_rounds(const K3MParams& q, const uint4 (&lr)[2], uint32_t t, uint32_t round) {
    const K3Params& p = q.k;
    const uint32_t off = round * 64;
    if (off >= p.pitch) return;
    const uint32_t bytes = min(256u, p.pitch - off);
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        if (lr[h].x == kLeanNo) continue;
        const uint32_t row = t == 0 ? lr[h].x : t == 1 ? lr[h].y : t == 2 ? lr[h].z : lr[h].w;
        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p.rows + size_t(row) * p.pitch + off), "r"(bytes)
                     : "memory");
    }
}

__device__ __forceinline__ uint32_t u4_word(const uint4& v, int j) { return j == 0 ? v.x : j == 1 ? v.y : j == 2 ? v.z : v.w; }

__device__ __forceinline__ void mma_compute_round(const K3MParams& q, MmaLaneSums& z, const MmaRound& d, uint32_t chunk,
                                                  const uint8_t* bfrag_round) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        uint32_t a[2], b[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            a[h] = u4_word(d.v[h][0], j) & u4_word(d.v[h][1], j);
            b[h] = u4_word(d.v[h][2], j) & u4_word(d.v[h][3], j);
        }
        mma_word_step(q, z, a, b, chunk * 4 + j, bfrag_round + j * 512);
    }
}

// tile epilogue shared by the two K3M kernels: the 4 lanes of a group hold the pieces of the same two snarls
__device__ __forceinline__ void mma_tile_epilogue(const K3MParams& q, MmaLaneSums& z, const uint4 (&lr)[2], uint64_t tile,
                                                  uint32_t slice, uint32_t g, uint32_t t, double sc0, double sc1) {
    const K3Params& p = q.k;
    double xs[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        xs[h] = double(z.c[2 * h]) * sc0 + double(z.c[2 * h + 1]) * sc1;
#pragma unroll
        for (int o = 1; o <= 2; o <<= 1) {
            z.cA[h] += __shfl_xor_sync(0xffffffffu, z.cA[h], o);
            z.cB[h] += __shfl_xor_sync(0xffffffffu, z.cB[h], o);
            z.cC[h] += __shfl_xor_sync(0xffffffffu, z.cC[h], o);
            z.nd[h] |= __shfl_xor_sync(0xffffffffu, z.nd[h], o);
            xs[h] += __shfl_xor_sync(0xffffffffu, xs[h], o);
        }
    }
    if (__any_sync(0xffffffffu, z.saw_unused)) {
#pragma unroll
        for (int h = 0; h < 2; ++h)
#pragma unroll
            for (int o = 1; o <= 2; o <<= 1) {
                z.su[h] += __shfl_xor_sync(0xffffffffu, z.su[h], o);
                z.suu[h] += __shfl_xor_sync(0xffffffffu, z.suu[h], o);
            }
    }
    if (t == 0) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const uint64_t s = tile * kMmaTile + g + 8u * h;
            if (s >= p.n_snarls) continue;
            if (z.live[h] && z.nd[h] != 0u && q.rec2x != nullptr) {
                // a sample on three path-haplotypes (overlapping paths): dosages of 1/3, 2/3 -> the 2x kernel
                mma_push_thirds(q.rec2x, q.count2x, q.count2x_dyn, p.mom, uint32_t(s), lr[h]);
            } else if (z.live[h]) {
                uint32_t* pi = q.part.i + size_t(slice) * 7 * q.part.stride + s;
                double* pd = q.part.d + size_t(slice) * 3 * q.part.stride + s;
                pi[0] = z.cA[h] & 0xFFFFu;
                pi[q.part.stride] = z.cA[h] >> 16;
                pi[2 * q.part.stride] = z.cB[h] & 0xFFFFu;
                pi[3 * q.part.stride] = z.cB[h] >> 16;
                pi[4 * q.part.stride] = z.cC[h];
                pi[5 * q.part.stride] = 0u;
                pi[6 * q.part.stride] = z.nd[h] ? kLeanReplayMark : 0u;  // K4 sends the snarl to the general path
                pd[0] = 0.5 * q.y_scale * xs[h];  // d = (2d) / 2
                pd[q.part.stride] = z.su[h];
                pd[2 * q.part.stride] = z.suu[h];
                if (slice == 0) p.mom[s].state = 2;  // K4 builds the moments from the partial sums
            } else if (slice == 0) {
                p.mom[s].state = 0;
                if (lr[h].y != 0u) {  // no paths at all: prep already put the snarl on the general list
                    const uint32_t slot = atomicAdd(q.slow_count, 1u);
                    q.slow_list[slot] = uint32_t(s);
                }
            }
        }
    }
}

// =================================================================================================
// K3M kernel.  grid = (CTAs, haplotype slices); every WARP owns tiles of 16 consecutive snarls (no barrier, no shared
// state between warps but the read-only slice table).  dynamic shared memory: the slice table of this haplotype slice.
// =================================================================================================
__global__ void __launch_bounds__(kMmaWarps * 32, 2) dosage_moments_mma_kernel(const K3MParams q) {
    extern __shared__ __align__(128) uint8_t ysl[];
    const K3Params& p = q.k;
    const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
    const uint32_t slice = blockIdx.y;
    const uint32_t r_first = slice * q.rounds_per_slice;
    const uint32_t r_len = min(q.rounds_per_slice, q.n_rounds - r_first);
    {
        const uint4* src = reinterpret_cast<const uint4*>(q.yslices + size_t(r_first) * 2048);
        uint4* dst = reinterpret_cast<uint4*>(ysl);
        for (uint32_t i = tid; i < r_len * 128u; i += blockDim.x) dst[i] = __ldg(src + i);
    }
    __syncthreads();
    const uint8_t* bfrag_lane = ysl + lane * 8;

    const uint64_t n_tiles = (p.n_snarls + kMmaTile - 1) / kMmaTile;
    const uint64_t stride = uint64_t(gridDim.x) * kMmaWarps;
    // column scales of the Horner sum: lane t owns the slices 2t, 2t + 1 of the accumulator rows: 128^-(col + 1)
    const double sc0 = __hiloint2double(int((1023u - 7u * (2u * t + 1u)) << 20), 0);
    const double sc1 = __hiloint2double(int((1023u - 7u * (2u * t + 2u)) << 20), 0);

    uint64_t tile = uint64_t(blockIdx.x) * kMmaWarps + warp;
    uint4 lr[2], lr_next[2];
    auto load_recs = [&](uint64_t tl, uint4 (&dst)[2]) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const uint64_t s = tl * kMmaTile + g + 8u * h;
            dst[h] = (tl < n_tiles && s < p.n_snarls) ? __ldg(q.lean_rec + s) : make_uint4(kLeanNo, 0u, 0u, 0u);
        }
    };
    load_recs(tile, lr);
    for (; tile < n_tiles; tile += stride) {
        load_recs(tile + stride, lr_next);
        MmaLaneSums z;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            z.cA[h] = z.cB[h] = z.cC[h] = z.nd[h] = 0u;
            z.su[h] = z.suu[h] = 0.0;
            z.live[h] = (lr[h].x != kLeanNo) ? 0xFFFFFFFFu : 0u;
        }
        z.c[0] = z.c[1] = z.c[2] = z.c[3] = 0;
        z.saw_unused = false;
        // ---- rounds: software pipelined two deep (A / B register sets) ------------------------------------------
        MmaRound ra, rb;
        mma_load_round(q, lr, z.live, (r_first + 0) * 4 + t, ra);
        if (q.prefetch) {
            mma_prefetch_rounds(q, lr, t, r_first + 4);       // rounds 4..7 of this tile
            mma_prefetch_rounds(q, lr_next, t, r_first + 0);  // rounds 0..3 of the warp's next tile
        }
        for (uint32_t c = 0; c < r_len; c += 2) {
            if (q.prefetch && (c & 3u) == 0u && c + 8 < r_len + 4) mma_prefetch_rounds(q, lr, t, r_first + c + 8);
            mma_load_round(q, lr, z.live, (c + 1 < r_len) ? (r_first + c + 1) * 4 + t : 0xFFFFFFF0u, rb);
            mma_compute_round(q, z, ra, (r_first + c) * 4 + t, bfrag_lane + size_t(c) * 2048);
            mma_load_round(q, lr, z.live, (c + 2 < r_len) ? (r_first + c + 2) * 4 + t : 0xFFFFFFF0u, ra);
            if (c + 1 < r_len) mma_compute_round(q, z, rb, (r_first + c + 1) * 4 + t, bfrag_lane + size_t(c + 1) * 2048);
        }
        mma_tile_epilogue(q, z, lr, tile, slice, g, t, sc0, sc1);
        lr[0] = lr_next[0];
        lr[1] = lr_next[1];
    }
}

// =================================================================================================
// K3M, ring variant (the default): the same tiles, class counts and MMAs, but the rows arrive through a per-warp
// shared-memory ring filled with cp.async instead of through registers.
//   * a copy instruction of the warp moves 128 contiguous bytes (a whole L2 line) of 4 rows (8 lanes x 16 B per row); a
//     stage is 128 B of each of the tile's 64 rows = 8 KB; S stages per warp.  Nothing in flight lives in registers, so
//     the loop holds one round (32 registers) instead of two rounds in flight + the one being worked on.
//   * within a stage row r = h * 32 + g * 4 + k of the tile sits at r * 128, its 16-byte chunk c at (c ^ 4 (g & 1)) * 16:
//     the quarter-warp of an LDS.128 (g in {2q, 2q + 1}, t = 0..3) then touches all 32 banks once.
//   * lane (g, t) reads chunks t and 4 + t of its 8 rows back as the two rounds 2m, 2m + 1 of macro step m: the chunk
//     numbering, the B fragments and the arithmetic are those of the register kernel, bit for bit.
// Measured on the bare access pattern (scripts/micro/k3m_pattern.cu, 5.1 GB, no arithmetic): 64 B x 8 rows through
// registers 5.4-6.1 TB/s with 16 warps per SM, this ring 6.7-6.8 TB/s with 8.
// =================================================================================================
constexpr int kRingStageBytes = 64 * 128;

template <int WARPS, int S>
__global__ void __launch_bounds__(WARPS * 32, 1) dosage_moments_mma_ring_kernel(const K3MParams q) {
    extern __shared__ __align__(128) uint8_t ysl[];
    const K3Params& p = q.k;
    const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
    const uint32_t slice = blockIdx.y;
    const uint32_t r_first = slice * q.rounds_per_slice;
    const uint32_t r_len = min(q.rounds_per_slice, q.n_rounds - r_first);
    {
        const uint4* src = reinterpret_cast<const uint4*>(q.yslices + size_t(r_first) * 2048);
        uint4* dst = reinterpret_cast<uint4*>(ysl);
        for (uint32_t i = tid; i < r_len * 128u; i += blockDim.x) dst[i] = __ldg(src + i);
    }
    __syncthreads();
    const uint8_t* bfrag_lane = ysl + lane * 8;
    uint8_t* ring = ysl + size_t(q.rounds_per_slice) * 2048 + size_t(warp) * S * kRingStageBytes;
    const uint32_t ring_s = uint32_t(__cvta_generic_to_shared(ring));
    // copy mapping: instruction i of a stage moves the rows of snarl (h = i >> 3, g = i & 7); this lane takes row k = sub
    const uint32_t sub = lane >> 3, c8 = lane & 7;
    // read mapping: this lane's 8 rows, chunk t (+ 4 for the odd round), swizzled by g & 1
    const uint32_t rd_chunk = (t ^ ((g & 1u) << 2)) * 16;

    const uint64_t n_tiles = (p.n_snarls + kMmaTile - 1) / kMmaTile;
    const uint64_t stride = uint64_t(gridDim.x) * WARPS;
    const double sc0 = __hiloint2double(int((1023u - 7u * (2u * t + 1u)) << 20), 0);
    const double sc1 = __hiloint2double(int((1023u - 7u * (2u * t + 2u)) << 20), 0);
    const uint32_t n_macro = (r_len + 1) / 2;
    const uint32_t chunk_end = (r_first + r_len) * 4 < p.pitch / 16 ? (r_first + r_len) * 4 : p.pitch / 16;

    uint64_t tile = uint64_t(blockIdx.x) * WARPS + warp;
    uint4 lr[2], lr_next[2];
    auto load_recs = [&](uint64_t tl, uint4 (&dst)[2]) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const uint64_t s = tl * kMmaTile + g + 8u * h;
            dst[h] = (tl < n_tiles && s < p.n_snarls) ? __ldg(q.lean_rec + s) : make_uint4(kLeanNo, 0u, 0u, 0u);
        }
    };
    load_recs(tile, lr);
    for (; tile < n_tiles; tile += stride) {
        load_recs(tile + stride, lr_next);
        MmaLaneSums z;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            z.cA[h] = z.cB[h] = z.cC[h] = z.nd[h] = 0u;
            z.su[h] = z.suu[h] = 0.0;
            z.live[h] = (lr[h].x != kLeanNo) ? 0xFFFFFFFFu : 0u;
        }
        z.c[0] = z.c[1] = z.c[2] = z.c[3] = 0;
        z.saw_unused = false;
        // rows this lane copies: component t of the record of the lane it is fetched from equals that lane's k = sub
        uint32_t rid[16];
        {
            uint32_t comp[2];
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const uint32_t v = t == 0 ? lr[h].x : t == 1 ? lr[h].y : t == 2 ? lr[h].z : lr[h].w;
                comp[h] = z.live[h] ? v : kLeanNo;
            }
#pragma unroll
            for (int i = 0; i < 16; ++i) rid[i] = __shfl_sync(0xffffffffu, comp[i >> 3], (i & 7) * 4 + sub);
        }
        auto issue = [&](uint32_t m) {
            if (m < n_macro) {
                const uint32_t chunk = (r_first + 2 * m) * 4 + c8;  // 16-byte chunk of the row this lane copies
                const bool in = chunk < chunk_end;
                const uint32_t dst0 = ring_s + (m % S) * kRingStageBytes + sub * 128 + c8 * 16;
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const bool ok = in && rid[i] != kLeanNo;
                    const uint8_t* src = p.rows + (ok ? size_t(rid[i]) * p.pitch + size_t(chunk) * 16 : size_t(0));
                    // row i * 4 + sub; g = i & 7 -> swizzle bit i & 1
                    const uint32_t dst = (dst0 + i * 512) ^ ((uint32_t(i) & 1u) << 6);
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(ok ? 16u : 0u)
                                 : "memory");
                }
            }
            asm volatile("cp.async.commit_group;" ::: "memory");
        };
#pragma unroll
        for (int m = 0; m < S - 1; ++m) issue(m);
        for (uint32_t m = 0; m < n_macro; ++m) {
            issue(m + S - 1);
            asm volatile("cp.async.wait_group %0;" ::"n"(S - 1) : "memory");
            __syncwarp();
            const uint8_t* st = ring + (m % S) * kRingStageBytes + g * 512;
#pragma unroll
            for (int sr = 0; sr < 2; ++sr) {
                const uint32_t round = 2 * m + sr;
                if (round < r_len) {
                    MmaRound d;
#pragma unroll
                    for (int h = 0; h < 2; ++h)
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            d.v[h][k] = *reinterpret_cast<const uint4*>(st + h * 4096 + k * 128 + (rd_chunk ^ (sr * 64u)));
                    mma_compute_round(q, z, d, (r_first + round) * 4 + t, bfrag_lane + size_t(round) * 2048);
                }
            }
            __syncwarp();
        }
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        mma_tile_epilogue(q, z, lr, tile, slice, g, t, sc0, sc1);
        lr[0] = lr_next[0];
        lr[1] = lr_next[1];
    }
}

// =================================================================================================
// K3M3: the same contraction for THREE-path snarls.  The regression keeps the columns of path 0 and path 1 (path 2 is
// the baseline the reference drops, snarl_analyzer.cpp:676-684), so the two M rows of a lane group (g, g + 8) are the
// two dosage columns of ONE snarl and a warp's tile is 8 snarls of the compacted three-path list (lean_rec_kernel).
// X^T X comes from popcounts of the class masks quant_fast<3> uses (kernels_quant.cu), X^T y from the integer MMA, and
// the lane that ends up with a snarl's sums applies the filters (snarl_analyzer.cpp:542-579) and writes its moments —
// no partial sums, so this kernel only runs when one CTA row covers the whole haplotype range (gridDim.y == 1).
// Data-dependent shapes the fast formulas do not cover (a sample on 3+ path-haplotypes, an empty path, two identical
// columns that the reference merges, :620-660) are pushed to the slow list: K3's warp-per-snarl kernel runs after this.
// =================================================================================================
constexpr int kMma3Tile = 8;

struct Mma3Sums {
    int c[4];
    uint32_t cAl01, cAl2U, cN1, cNh, cN2, cHh;  // packed pairs: allele 0 | 1, allele 2 | n_used, n1, nh, n2 of col 0 | 1
    uint32_t d01, d02, d12, bad;               // OR of the differences between columns; samples with row sum >= 3
    double su, suu;                            // sum y, sum y^2 over the samples on no path
    uint32_t live;
    bool saw_unused;
};

__device__ __forceinline__ void bitslice_add(uint32_t& s0, uint32_t& s1, uint32_t& ovf, uint32_t x) {
    const uint32_t c = s0 & x;
    s0 ^= x;
    ovf |= s1 & c;
    s1 ^= c;
}

__device__ __forceinline__ void mma3_word_step(const K3MParams& q, Mma3Sums& z, uint32_t a, uint32_t b, uint32_t c,
                                               uint32_t word, const uint8_t* bfrag) {
    const uint32_t M = 0x55555555u;
    const uint2 bm0 = *reinterpret_cast<const uint2*>(bfrag);
    const uint2 bm1 = *reinterpret_cast<const uint2*>(bfrag + 256);
    const uint32_t lo0 = a & M, hi0 = (a >> 1) & M, lo1 = b & M, hi1 = (b >> 1) & M, lo2 = c & M, hi2 = (c >> 1) & M;
    const uint32_t c1_0 = lo0 ^ hi0, c2_0 = lo0 & hi0, c1_1 = lo1 ^ hi1, c2_1 = lo1 & hi1, c1_2 = lo2 ^ hi2, c2_2 = lo2 & hi2;
    uint32_t s0 = lo0, s1 = 0u, ovf = 0u;  // bit-sliced row sum of the six presence bits
    bitslice_add(s0, s1, ovf, hi0);
    bitslice_add(s0, s1, ovf, lo1);
    bitslice_add(s0, s1, ovf, hi1);
    bitslice_add(s0, s1, ovf, lo2);
    bitslice_add(s0, s1, ovf, hi2);
    const uint32_t a1 = s0 & ~s1, a2 = s1 & ~s0, used = s0 | s1 | ovf;
    z.bad |= ovf | (s0 & s1);
    z.cAl01 += __popc(a) | (__popc(b) << 16);
    z.cAl2U += __popc(c) | (__popc(used) << 16);
    z.cN1 += __popc(c1_0 & a1) | (__popc(c1_1 & a1) << 16);
    z.cNh += __popc(c1_0 & a2) | (__popc(c1_1 & a2) << 16);
    z.cN2 += __popc(c2_0) | (__popc(c2_1) << 16);
    z.cHh += __popc(c1_0 & c1_1 & a2);
    z.d01 |= (c1_0 ^ c1_1) | (c2_0 ^ c2_1);
    z.d02 |= (c1_0 ^ c1_2) | (c2_0 ^ c2_2);
    z.d12 |= (c1_1 ^ c1_2) | (c2_1 ^ c2_2);
    // 2-bit fields 2 * dosage of the two regression columns: 1 = half (one of two), 2 = the only / both haplotypes
    const uint32_t f0 = (c1_0 & a2) | (((c1_0 & a1) | c2_0) << 1);
    const uint32_t f1 = (c1_1 & a2) | (((c1_1 & a1) | c2_1) << 1);
    mma_s8_16832(z.c, f0 & 0x03030303u, f1 & 0x03030303u, (f0 >> 2) & 0x03030303u, (f1 >> 2) & 0x03030303u, bm0.x, bm0.y);
    mma_s8_16832(z.c, (f0 >> 4) & 0x03030303u, (f1 >> 4) & 0x03030303u, (f0 >> 6) & 0x03030303u, (f1 >> 6) & 0x03030303u,
                 bm1.x, bm1.y);
    uint32_t un = ~used & M & valid_word(word, q.k.n_hap) & z.live;
    if (un != 0u) {
        z.saw_unused = true;
        while (un) {
            const int bit = __ffs(un) - 1;
            un &= un - 1;
            const double yv = __ldg(q.yc + size_t(word) * 16 + (bit >> 1));
            z.su += yv;
            z.suu = fma(yv, yv, z.suu);
        }
    }
}

struct Mma3Round {
    uint4 v[6];
};
__device__ __forceinline__ void mma3_load_round(const K3MParams& q, const uint4& ra, const uint4& rb, uint32_t live,
                                                uint32_t chunk, Mma3Round& d) {
    const K3Params& p = q.k;
    const bool in = chunk < p.pitch / 16 && live != 0u;
    const uint32_t rr[6] = {ra.x, ra.y, ra.z, ra.w, rb.x, rb.y};
#pragma unroll
    for (int k = 0; k < 6; ++k) {
        d.v[k] = make_uint4(0u, 0u, 0u, 0u);
        if (in) d.v[k] = ldg_stream(reinterpret_cast<const uint4*>(p.rows + size_t(rr[k]) * p.pitch + size_t(chunk) * 16));
    }
}
__device__ __forceinline__ void mma3_compute_round(const K3MParams& q, Mma3Sums& z, const Mma3Round& d, uint32_t chunk,
                                                   const uint8_t* bfrag_round) {
#pragma unroll
    for (int j = 0; j < 4; ++j)
        mma3_word_step(q, z, u4_word(d.v[0], j) & u4_word(d.v[1], j), u4_word(d.v[2], j) & u4_word(d.v[3], j),
                       u4_word(d.v[4], j) & u4_word(d.v[5], j), chunk * 4 + j, bfrag_round + j * 512);
}

__global__ void __launch_bounds__(kMmaWarps * 32, 2) dosage_moments_mma3_kernel(const K3MParams q,
                                                                                const uint4* __restrict__ rec3,
                                                                                const uint32_t* __restrict__ count3) {
    extern __shared__ __align__(128) uint8_t ysl[];
    const K3Params& p = q.k;
    const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
    const uint32_t n3 = *count3;
    if (n3 == 0u) return;
    const uint32_t r_len = q.n_rounds;
    if (!q.table_in_global) {
        const uint4* src = reinterpret_cast<const uint4*>(q.yslices);
        uint4* dst = reinterpret_cast<uint4*>(ysl);
        for (uint32_t i = tid; i < r_len * 128u; i += blockDim.x) dst[i] = __ldg(src + i);
    }
    __syncthreads();
    const uint8_t* bfrag_lane = (q.table_in_global ? q.yslices : ysl) + lane * 8;
    const uint32_t n_tiles = (n3 + kMma3Tile - 1) / kMma3Tile;
    const uint32_t stride = gridDim.x * kMmaWarps;
    const double sc0 = __hiloint2double(int((1023u - 7u * (2u * t + 1u)) << 20), 0);
    const double sc1 = __hiloint2double(int((1023u - 7u * (2u * t + 2u)) << 20), 0);

    for (uint32_t tile = blockIdx.x * kMmaWarps + warp; tile < n_tiles; tile += stride) {
        const uint32_t idx = tile * kMma3Tile + g;
        uint4 ra = make_uint4(0u, 0u, 0u, 0u), rb = ra;
        Mma3Sums z;
        z.live = 0u;
        if (idx < n3) {
            ra = __ldg(rec3 + 2 * size_t(idx));
            rb = __ldg(rec3 + 2 * size_t(idx) + 1);
            z.live = 0xFFFFFFFFu;
        }
        z.c[0] = z.c[1] = z.c[2] = z.c[3] = 0;
        z.cAl01 = z.cAl2U = z.cN1 = z.cNh = z.cN2 = z.cHh = 0u;
        z.d01 = z.d02 = z.d12 = z.bad = 0u;
        z.su = z.suu = 0.0;
        z.saw_unused = false;
        Mma3Round da, db;
        mma3_load_round(q, ra, rb, z.live, t, da);
        for (uint32_t c = 0; c < r_len; c += 2) {
            mma3_load_round(q, ra, rb, z.live, (c + 1 < r_len) ? (c + 1) * 4 + t : 0xFFFFFFF0u, db);
            mma3_compute_round(q, z, da, c * 4 + t, bfrag_lane + size_t(c) * 2048);
            mma3_load_round(q, ra, rb, z.live, (c + 2 < r_len) ? (c + 2) * 4 + t : 0xFFFFFFF0u, da);
            if (c + 1 < r_len) mma3_compute_round(q, z, db, (c + 1) * 4 + t, bfrag_lane + size_t(c + 1) * 2048);
        }
        // ---- the 4 lanes of a group hold the pieces of one snarl ---------------------------------------------------
        double xs[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) xs[h] = double(z.c[2 * h]) * sc0 + double(z.c[2 * h + 1]) * sc1;
#pragma unroll
        for (int o = 1; o <= 2; o <<= 1) {
            z.cAl01 += __shfl_xor_sync(0xffffffffu, z.cAl01, o);
            z.cAl2U += __shfl_xor_sync(0xffffffffu, z.cAl2U, o);
            z.cN1 += __shfl_xor_sync(0xffffffffu, z.cN1, o);
            z.cNh += __shfl_xor_sync(0xffffffffu, z.cNh, o);
            z.cN2 += __shfl_xor_sync(0xffffffffu, z.cN2, o);
            z.cHh += __shfl_xor_sync(0xffffffffu, z.cHh, o);
            z.d01 |= __shfl_xor_sync(0xffffffffu, z.d01, o);
            z.d02 |= __shfl_xor_sync(0xffffffffu, z.d02, o);
            z.d12 |= __shfl_xor_sync(0xffffffffu, z.d12, o);
            z.bad |= __shfl_xor_sync(0xffffffffu, z.bad, o);
            xs[0] += __shfl_xor_sync(0xffffffffu, xs[0], o);
            xs[1] += __shfl_xor_sync(0xffffffffu, xs[1], o);
        }
        if (__any_sync(0xffffffffu, z.saw_unused)) {
#pragma unroll
            for (int o = 1; o <= 2; o <<= 1) {
                z.su += __shfl_xor_sync(0xffffffffu, z.su, o);
                z.suu += __shfl_xor_sync(0xffffffffu, z.suu, o);
            }
        }
        if (t != 0 || !z.live) continue;
        const uint64_t s = rb.w;
        const uint32_t p0 = rb.z;
        const uint32_t al0 = z.cAl01 & 0xFFFFu, al1 = z.cAl01 >> 16, al2 = z.cAl2U & 0xFFFFu, n_used = z.cAl2U >> 16;
        p.out.allele_paths[p0] = al0;  // quantitative_table.cpp:54-55
        p.out.allele_paths[p0 + 1] = al1;
        p.out.allele_paths[p0 + 2] = al2;
        if (z.bad != 0u || al0 == 0u || al1 == 0u || al2 == 0u || z.d01 == 0u || z.d02 == 0u || z.d12 == 0u) {
            p.mom[s].state = 0;
            const uint32_t slot = atomicAdd(q.slow_count, 1u);
            q.slow_list[slot] = uint32_t(s);
            continue;
        }
        const double n1_0 = double(z.cN1 & 0xFFFFu), n1_1 = double(z.cN1 >> 16);
        const double nh_0 = double(z.cNh & 0xFFFFu), nh_1 = double(z.cNh >> 16);
        const double n2_0 = double(z.cN2 & 0xFFFFu), n2_1 = double(z.cN2 >> 16);
        const double total = double(n_used);  // every used row sums to 1 (no sample with row sum >= 3)
        const double col0 = n1_0 + 0.5 * nh_0 + n2_0, col1 = n1_1 + 0.5 * nh_1 + n2_1, col2 = total - col0 - col1;
        bool filtered = uint64_t(n_used) < p.thr.min_individuals;
        if (!filtered) {
            if (total < double(p.thr.min_haplotypes)) filtered = true;
            const double cols[3] = {col0, col1, col2};
            int above = 0;
#pragma unroll
            for (int j = 0; j < 3; ++j) {
                const double freq = __ddiv_rn(cols[j], total);
                const double maf = fmin(freq, __dsub_rn(1.0, freq));
                if (maf > p.thr.maf_threshold) ++above;
            }
            if (above < 2) filtered = true;
        }
        p.out.filtered[s] = filtered ? 1 : 0;
        p.out.n_used[s] = n_used;
        if (filtered) {
            p.mom[s].state = 0;
            continue;
        }
        QuantMoments mom;
#pragma unroll
        for (int i = 0; i < 10; ++i) mom.xtx[i] = 0.0;
        mom.xty[3] = 0.0;
        const double su = p.sy_all - z.su, suu = p.syy_all - z.suu;
        mom.n_used = n_used;
        mom.p = 3;
        mom.state = 1;
        mom.pad = 0;
        mom.xtx[0] = total;
        mom.xtx[1] = col0;
        mom.xtx[2] = n1_0 + 0.25 * nh_0 + n2_0;
        mom.xtx[3] = col1;
        mom.xtx[4] = 0.25 * double(z.cHh);
        mom.xtx[5] = n1_1 + 0.25 * nh_1 + n2_1;
        mom.xty[0] = su;
        mom.xty[1] = 0.5 * q.y_scale * xs[0];
        mom.xty[2] = 0.5 * q.y_scale * xs[1];
        mom.sy = su;
        mom.syy = suu;
        p.mom[s] = mom;
    }
}

// =================================================================================================
// K3M2x: two-path snarls with NON-DYADIC dosages — a sample on three path-haplotypes has d0 = 1/3 or 2/3.  That is every
// carrier of the other path when one path has no checked edge ('*', it matches everybody: known from the path program,
// lean_rec_kernel lists those), and the occasional sample of a snarl with overlapping paths (K3M meets them and lists
// the snarl).  The two M rows of a lane group are the two halves of ONE snarl's dosage column: 2d over the dyadic
// classes (half, full) and 3d over the thirds (third, two thirds), so  sum d y = scale * (C_A / 2 + C_B / 3)  with both
// contractions exact integers.  Class counts as in quant_lean2; filters and moments by lean_finish.  One haplotype slice
// only (like K3M3).
// =================================================================================================
struct Mma2xSums {
    int c[4];
    uint32_t cAl, cUF, cHT, cTT;  // al0 | al1, n_used | n_full0, n_half | n_third, n_tthird
    double su, suu;
    uint32_t live;
    bool saw_unused;
};

__device__ __forceinline__ void mma2x_word_step(const K3MParams& q, Mma2xSums& z, uint32_t a, uint32_t b, uint32_t word,
                                                const uint8_t* bfrag) {
    const uint2 bm0 = *reinterpret_cast<const uint2*>(bfrag);
    const uint2 bm1 = *reinterpret_cast<const uint2*>(bfrag + 256);
    const LeanMasks m = lean_masks(a, b);
    z.cAl += __popc(a) | (__popc(b) << 16);
    z.cUF += __popc(m.used) | (__popc(m.full0) << 16);
    z.cHT += __popc(m.half) | (__popc(m.third) << 16);
    z.cTT += __popc(m.tthird);
    const uint32_t f0 = m.half | (m.full0 << 1);    // 2 * d0 where it is 1/2 or 1
    const uint32_t f1 = m.third | (m.tthird << 1);  // 3 * d0 where it is 1/3 or 2/3
    mma_s8_16832(z.c, f0 & 0x03030303u, f1 & 0x03030303u, (f0 >> 2) & 0x03030303u, (f1 >> 2) & 0x03030303u, bm0.x, bm0.y);
    mma_s8_16832(z.c, (f0 >> 4) & 0x03030303u, (f1 >> 4) & 0x03030303u, (f0 >> 6) & 0x03030303u, (f1 >> 6) & 0x03030303u,
                 bm1.x, bm1.y);
    uint32_t un = ~m.used & 0x55555555u & valid_word(word, q.k.n_hap) & z.live;
    if (un != 0u) {
        z.saw_unused = true;
        while (un) {
            const int bit = __ffs(un) - 1;
            un &= un - 1;
            const double yv = __ldg(q.yc + size_t(word) * 16 + (bit >> 1));
            z.su += yv;
            z.suu = fma(yv, yv, z.suu);
        }
    }
}

// The filters of a two-path snarl whose decision lies within rounding of a threshold (lean_finish returned 2): exact
// replay of filtration_quantitative_table's accumulation order (src/snarl_analyzer.cpp:558-576) — `total` is ONE running
// sum over (sample, column), the column sums run down the samples — with the table entries count / row sum of
// quantitative_table.cpp:93.  With thirds in the table the outcome of `maf > threshold` at an exact tie (250 carriers
// x 1/3 against 0.05 x 5000 ...) depends on how those 5000 additions round.  One lane, serial by definition: three
// independent chains of FP64 additions, ~0.1 ms — the general path took 1.4 ms for the same replay.
struct ExactArgs {
    const uint8_t* rows;
    uint32_t pitch, n_hap;
    double min_haplotypes, maf_threshold;
    const double2* tab;  // [16] (d0, d1) by the presence bits of a sample: path 0 bits | path 1 bits << 2
};
// table entry count / row sum of quantitative_table.cpp:93 for the sample with presence bits `idx`
__device__ __forceinline__ double2 exact_tab_entry(uint32_t idx) {
    const uint32_t k0 = __popc(idx & 3u), k1 = __popc(idx >> 2), rs = k0 + k1;
    double2 d;
    d.x = k0 ? __ddiv_rn(double(k0), double(rs)) : 0.0;
    d.y = k1 ? __ddiv_rn(double(k1), double(rs)) : 0.0;
    return d;
}
__device__ __noinline__ bool lean_exact_filtered(const ExactArgs p, uint4 rows) {
    const uint32_t rr[4] = {rows.x, rows.y, rows.z, rows.w};
    const uint32_t n_chunks = (p.n_hap + 127u) / 128u;  // <= pitch / 16
    double cs0 = 0.0, cs1 = 0.0, tot = 0.0;
    auto load = [&](uint32_t c, uint4 (&v)[4]) {
#pragma unroll
        for (int k = 0; k < 4; ++k)
            v[k] = rr[k] == kLeanAll
                       ? make_uint4(valid_word(4 * c, p.n_hap), valid_word(4 * c + 1, p.n_hap),
                                    valid_word(4 * c + 2, p.n_hap), valid_word(4 * c + 3, p.n_hap))
                       : __ldg(reinterpret_cast<const uint4*>(p.rows + size_t(rr[k]) * p.pitch) + c);
    };
    uint4 cur[4], nxt[4];
    load(0, cur);
    for (uint32_t c = 0; c < n_chunks; ++c) {
        if (c + 1 < n_chunks) load(c + 1, nxt);  // in flight while this chunk's 64 samples are added
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const uint32_t a = u4_word(cur[0], j) & u4_word(cur[1], j), b = u4_word(cur[2], j) & u4_word(cur[3], j);
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                // a sample on no path has the entry (0, 0): adding +0.0 to a non-negative sum changes nothing
                const double2 d = p.tab[((a >> (2 * i)) & 3u) | (((b >> (2 * i)) & 3u) << 2)];
                cs0 = __dadd_rn(cs0, d.x);
                cs1 = __dadd_rn(cs1, d.y);
                tot = __dadd_rn(__dadd_rn(tot, d.x), d.y);
            }
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) cur[k] = nxt[k];
    }
    bool filt = tot < p.min_haplotypes;  // :565
    int above = 0;
    const double cs[2] = {cs0, cs1};
#pragma unroll
    for (int c = 0; c < 2; ++c) {  // :570-576
        const double freq = __ddiv_rn(cs[c], tot);
        const double maf = fmin(freq, __dsub_rn(1.0, freq));
        if (maf > p.maf_threshold) ++above;
    }
    return filt || above < 2;
}

struct Mma2xRound {
    uint4 v[4];
};
__device__ __forceinline__ void mma2x_load_round(const K3MParams& q, const uint4& ra, uint32_t live, uint32_t chunk,
                                                 Mma2xRound& d) {
    const K3Params& p = q.k;
    const bool in = chunk < p.pitch / 16 && live != 0u;
    const uint32_t rr[4] = {ra.x, ra.y, ra.z, ra.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        d.v[k] = make_uint4(0u, 0u, 0u, 0u);
        if (in) {
            if (rr[k] == kLeanAll)
                d.v[k] = make_uint4(valid_word(4 * chunk, p.n_hap), valid_word(4 * chunk + 1, p.n_hap),
                                    valid_word(4 * chunk + 2, p.n_hap), valid_word(4 * chunk + 3, p.n_hap));
            else
                d.v[k] = ldg_stream(reinterpret_cast<const uint4*>(p.rows + size_t(rr[k]) * p.pitch + size_t(chunk) * 16));
        }
    }
}
__device__ __forceinline__ void mma2x_compute_round(const K3MParams& q, Mma2xSums& z, const Mma2xRound& d, uint32_t chunk,
                                                    const uint8_t* bfrag_round) {
#pragma unroll
    for (int j = 0; j < 4; ++j)
        mma2x_word_step(q, z, u4_word(d.v[0], j) & u4_word(d.v[1], j), u4_word(d.v[2], j) & u4_word(d.v[3], j),
                        chunk * 4 + j, bfrag_round + j * 512);
}

__global__ void __launch_bounds__(kMmaWarps * 32, 2) dosage_moments_mma2x_kernel(const K3MParams q) {
    extern __shared__ __align__(128) uint8_t ysl[];
    const K3Params& p = q.k;
    const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
    const uint32_t n2 = *q.count2x + *q.count2x_dyn;
    if (n2 == 0u) return;
    __shared__ double2 exact_tab[16];  // lean_exact_filtered
    if (tid < 16) exact_tab[tid] = exact_tab_entry(tid);
    const uint32_t r_len = q.n_rounds;
    if (!q.table_in_global) {
        const uint4* src = reinterpret_cast<const uint4*>(q.yslices);
        uint4* dst = reinterpret_cast<uint4*>(ysl);
        for (uint32_t i = tid; i < r_len * 128u; i += blockDim.x) dst[i] = __ldg(src + i);
    }
    __syncthreads();
    const uint8_t* bfrag_lane = (q.table_in_global ? q.yslices : ysl) + lane * 8;
    const uint32_t n_tiles = (n2 + kMma3Tile - 1) / kMma3Tile;
    const uint32_t stride = gridDim.x * kMmaWarps;
    const double sc0 = __hiloint2double(int((1023u - 7u * (2u * t + 1u)) << 20), 0);
    const double sc1 = __hiloint2double(int((1023u - 7u * (2u * t + 2u)) << 20), 0);

    for (uint32_t tile = blockIdx.x * kMmaWarps + warp; tile < n_tiles; tile += stride) {
        const uint32_t idx = tile * kMma3Tile + g;
        uint4 ra = make_uint4(0u, 0u, 0u, 0u), rb = ra;
        Mma2xSums z;
        z.live = 0u;
        if (idx < n2) {
            ra = __ldcg(q.rec2x + 2 * size_t(idx));  // written by the kernels before this one
            rb = __ldcg(q.rec2x + 2 * size_t(idx) + 1);
            z.live = 0xFFFFFFFFu;
        }
        z.c[0] = z.c[1] = z.c[2] = z.c[3] = 0;
        z.cAl = z.cUF = z.cHT = z.cTT = 0u;
        z.su = z.suu = 0.0;
        z.saw_unused = false;
        Mma2xRound da, db;
        mma2x_load_round(q, ra, z.live, t, da);
        for (uint32_t c = 0; c < r_len; c += 2) {
            mma2x_load_round(q, ra, z.live, (c + 1 < r_len) ? (c + 1) * 4 + t : 0xFFFFFFF0u, db);
            mma2x_compute_round(q, z, da, c * 4 + t, bfrag_lane + size_t(c) * 2048);
            mma2x_load_round(q, ra, z.live, (c + 2 < r_len) ? (c + 2) * 4 + t : 0xFFFFFFF0u, da);
            if (c + 1 < r_len) mma2x_compute_round(q, z, db, (c + 1) * 4 + t, bfrag_lane + size_t(c + 1) * 2048);
        }
        double xs[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) xs[h] = double(z.c[2 * h]) * sc0 + double(z.c[2 * h + 1]) * sc1;
#pragma unroll
        for (int o = 1; o <= 2; o <<= 1) {
            z.cAl += __shfl_xor_sync(0xffffffffu, z.cAl, o);
            z.cUF += __shfl_xor_sync(0xffffffffu, z.cUF, o);
            z.cHT += __shfl_xor_sync(0xffffffffu, z.cHT, o);
            z.cTT += __shfl_xor_sync(0xffffffffu, z.cTT, o);
            xs[0] += __shfl_xor_sync(0xffffffffu, xs[0], o);
            xs[1] += __shfl_xor_sync(0xffffffffu, xs[1], o);
        }
        if (__any_sync(0xffffffffu, z.saw_unused)) {
#pragma unroll
            for (int o = 1; o <= 2; o <<= 1) {
                z.su += __shfl_xor_sync(0xffffffffu, z.su, o);
                z.suu += __shfl_xor_sync(0xffffffffu, z.suu, o);
            }
        }
        if (t != 0 || !z.live) continue;
        const uint64_t s = rb.w;
        const uint32_t p0 = *reinterpret_cast<const uint32_t*>(p.blob + size_t(p.rec[s].blob16) * 16 + 4);
        LeanSums ls;
        ls.al0 = z.cAl & 0xFFFFu;
        ls.al1 = z.cAl >> 16;
        ls.n_used = z.cUF & 0xFFFFu;
        ls.n_full0 = z.cUF >> 16;
        ls.n_half = z.cHT & 0xFFFFu;
        ls.n_third = z.cHT >> 16;
        ls.n_tthird = z.cTT;
        ls.xty1 = q.y_scale * (0.5 * xs[0] + (1.0 / 3.0) * xs[1]);
        ls.su_un = z.su;
        ls.suu_un = z.suu;
        QuantMoments mom;
        int r = lean_finish(p, s, p0, ls, mom);
        if (r == 2)  // a filter decision within rounding of its threshold: replay the reference's sums
            r = lean_finish(p, s, p0, ls, mom,
                            lean_exact_filtered(ExactArgs{p.rows, p.pitch, p.n_hap, double(p.thr.min_haplotypes),
                                                          p.thr.maf_threshold, exact_tab}, ra) ? 1 : 0);
        if (r == 1) p.mom[s] = mom;
        else p.mom[s].state = 0;
    }
}
