// kernels_binary.cu — the binary-trait hot path on sm_100a.
//
//   prep            per snarl: distinct rows + a compact "path program" (which of the snarl's rows each
//                   path ANDs), so that K1 touches every matrix row of a snarl exactly once.
//   K1 path_and_popc  identify_path (src/snarl_analyzer.cpp:315-356) + create_binary_table
//                   (src/binary_table.cpp:19-44): per path AND over its rows, popcount split by the
//                   case mask.  HBM-bound.  Rows are staged into shared memory by TMA bulk copies
//                   (cp.async.bulk + mbarrier), each warp running its own multi-stage pipeline.
//   K1' (LDG)       same arithmetic straight from global memory for snarls that do not fit a stage
//                   (many distinct rows / long path programs).
//   K2 binary_stats remove_empty_columns_binary_table + filtration_binary_table
//                   (src/snarl_analyzer.cpp:686-732) + FisherKhi2::fisher_khi2
//                   (src/stats_test.cpp:181-380), one thread per snarl, FP64.
//
// Path program ("blob") of a snarl, bytes:  [np] [3 reserved] [first path index, u32] [code_0 ..
// code_{np-1}] [locals of path 0] [locals of path 1] ...   code = 0xFF: an edge of the path is missing from the matrix => nobody matches
// (:334-336); code = 0: no checked edge => every haplotype matches (:343-354); else the number of
// checked edges R, followed (in the locals area) by R indices into the snarl's distinct-row list.
#include <cstdlib>

#include "ctx.hpp"
#include "device_math.cuh"
#include "ptx.cuh"
#include "stage_pipeline.cuh"

namespace stoat {

constexpr uint32_t kMissing = 0xFFFFFFFFu;
constexpr int kBlobBytes = kMaxBlob16 * 16;
constexpr uint32_t kBlobHeader = 8;

// =================================================================================================
// prep
// =================================================================================================
__global__ void __launch_bounds__(128) binary_prep_kernel(BatchDev b, uint64_t n_matrix_rows) {
    const uint64_t s = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (s >= b.n_snarls) return;
    const uint64_t p0 = b.snarl_path_off[s], p1 = b.snarl_path_off[s + 1];
    const uint64_t e0 = b.path_edge_off[p0], e1 = b.path_edge_off[p1];
    const uint64_t np = p1 - p0, ne = e1 - e0;
    const uint64_t base16 = 2 * s + ((p0 + e0) >> 4);  // room: 2 + floor((np+ne)/16) >= ceil((8+np+ne)/16)

    SnarlRec rec;
    for (int k = 0; k < kMaxStageRows; ++k) rec.rows[k] = 0;
    rec.blob16 = uint32_t(base16);
    rec.n_rows = 0;
    rec.blob_len16 = 1;
    rec.n_paths = 0;

    bool big = (np > 250) || (kBlobHeader + np + ne > uint64_t(kBlobBytes));
    uint32_t urows[kMaxStageRows];
    int nu = 0;
    // distinct rows (and validation of the row indices)
    for (uint64_t e = e0; e < e1; ++e) {
        const uint32_t r = b.edge_row[e];
        if (r == kMissing) continue;
        if (uint64_t(r) >= n_matrix_rows) {
            atomicExch(&b.counters[1], 1u);  // invalid row index: reported by the ABI call
            big = true;
            continue;
        }
        if (big) continue;
        int f = -1;
        for (int k = 0; k < nu; ++k)
            if (urows[k] == r) f = k;
        if (f < 0) {
            if (nu == kMaxStageRows)
                big = true;
            else
                urows[nu++] = r;
        }
    }
    uint8_t* blob = b.blob + base16 * 16;
    if (big) {
        // "nothing to do here" for the TMA kernel: the WHOLE header, not just np — the buffer is reused between batches
        // and a stale shape class (blob[1] == 1) would send the kernel down the biallelic branch with a stale p0
        *reinterpret_cast<uint2*>(blob) = make_uint2(0u, 0u);
        const uint32_t slot = atomicAdd(&b.counters[0], 1u);
        b.big_list[slot] = uint32_t(s);
        b.rec[s] = rec;
        return;
    }
    blob[0] = uint8_t(np);
    blob[1] = blob[2] = blob[3] = 0;  // blob[1]: shape class, set below
    *reinterpret_cast<uint32_t*>(blob + 4) = uint32_t(p0);
    uint32_t off = kBlobHeader + uint32_t(np);
    for (uint64_t j = 0; j < np; ++j) {
        const uint64_t a = b.path_edge_off[p0 + j], z = b.path_edge_off[p0 + j + 1];
        bool missing = false;
        for (uint64_t e = a; e < z; ++e) missing |= (b.edge_row[e] == kMissing);
        if (missing) {
            blob[kBlobHeader + j] = 0xFF;
            continue;
        }
        blob[kBlobHeader + j] = uint8_t(z - a);
        for (uint64_t e = a; e < z; ++e) {
            const uint32_t r = b.edge_row[e];
            int f = 0;
            for (int k = 0; k < nu; ++k)
                if (urows[k] == r) f = k;
            blob[off++] = uint8_t(f);
        }
    }
    // shape class 1: two paths with two checked edges each (K1's straight-line case)
    if (np == 2 && blob[kBlobHeader] == 2 && blob[kBlobHeader + 1] == 2) blob[1] = 1;
    for (int k = 0; k < nu; ++k) rec.rows[k] = urows[k];
    rec.n_rows = uint8_t(nu);
    rec.blob_len16 = uint8_t((off + 15) / 16);
    rec.n_paths = uint16_t(np);
    atomicMax(&b.counters[2], uint32_t(nu));  // widest stage needed
    b.rec[s] = rec;
}

cudaError_t launch_binary_prep(const BatchDev& b, uint64_t n_matrix_rows, cudaStream_t st) {
    if (b.n_snarls == 0) return cudaSuccess;
    const unsigned blocks = unsigned((b.n_snarls + 127) / 128);
    binary_prep_kernel<<<blocks, 128, 0, st>>>(b, n_matrix_rows);
    return cudaGetLastError();
}

// =================================================================================================
// K1: TMA-staged AND + popcount
// =================================================================================================
struct K1Params {
    const uint8_t* rows;       // repacked matrix
    const uint8_t* case_mask;  // pitch bytes
    const SnarlRec* rec;
    const uint8_t* blob;
    uint32_t* g0;  // [P]
    uint32_t* g1;  // [P]
    uint64_t n_snarls;
    uint32_t pitch;        // bytes per row
    uint32_t n_hap;        // n_cols
    uint32_t n_case_hap;   // popcount of the case mask
    uint32_t stage_rows;   // row slots per stage
    uint32_t n_stages;     // pipeline depth per warp
};

// AND of R staged rows for one 16-byte chunk, then the two popcounts, packed as all | case << 16 (n_hap < 65536) or
// kept apart.  R is a compile-time constant so the loads and ANDs are straight-line code.
template <int R>
__device__ __forceinline__ void path_counts(const uint8_t* rows_s, const uint8_t* mask_s, const uint32_t (&roff)[4],
                                            uint32_t lane, uint32_t chunks, uint32_t& n_all, uint32_t& n_case) {
    for (uint32_t c = lane; c < chunks; c += 32) {
        uint4 v = *reinterpret_cast<const uint4*>(rows_s + roff[0] + c * 16);
#pragma unroll
        for (int r = 1; r < R; ++r) v = and4(v, *reinterpret_cast<const uint4*>(rows_s + roff[r] + c * 16));
        const uint4 m = *reinterpret_cast<const uint4*>(mask_s + c * 16);
        n_all += popc4(v);
        n_case += popc4(and4(v, m));
    }
}

__global__ void __launch_bounds__(512, 1) path_and_popc_tma_kernel(const K1Params p) {
    extern __shared__ __align__(128) uint8_t smem[];
    const uint32_t n_warps = blockDim.x >> 5;
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t stage_bytes = kBlobBytes + p.stage_rows * p.pitch;
    // layout: [full barriers: n_warps * n_stages * 8][pad to 128][case mask: pitch][pad to 128][stages]
    uint64_t* full = reinterpret_cast<uint64_t*>(smem);
    const uint32_t bar_bytes = (n_warps * p.n_stages * 8 + 127) & ~127u;
    uint8_t* mask_s = smem + bar_bytes;
    const uint32_t mask_bytes = (p.pitch + 127) & ~127u;
    uint8_t* stages = mask_s + mask_bytes;

    if (threadIdx.x == 0) {
        for (uint32_t i = 0; i < n_warps * p.n_stages; ++i) mbar_init(&full[i], 1);
        mbar_fence_init();
    }
    for (uint32_t i = threadIdx.x; i < p.pitch / 16; i += blockDim.x)
        reinterpret_cast<uint4*>(mask_s)[i] = __ldg(reinterpret_cast<const uint4*>(p.case_mask) + i);
    __syncthreads();

    const uint64_t gw = uint64_t(blockIdx.x) * n_warps + warp;  // adjacent warps take adjacent snarls
    const uint64_t stride = uint64_t(gridDim.x) * n_warps;
    uint64_t* my_full = full + warp * p.n_stages;
    uint8_t* my_stages = stages + size_t(warp) * p.n_stages * stage_bytes;
    const uint32_t chunks = p.pitch / 16;
    const uint64_t policy = make_evict_first_policy();
    const bool packable = p.n_hap < 65536u;  // both counts of a path fit one 32-bit warp reduction

    // prologue: fill the pipeline
    {
        uint64_t s = gw;
        for (uint32_t st = 0; st < p.n_stages && s < p.n_snarls; ++st, s += stride)
            issue_stage(load_rec_word(p.rec, s, lane), lane, my_stages + size_t(st) * stage_bytes, &my_full[st], p.rows,
                        p.blob, p.pitch, policy);
    }

    uint32_t it = 0;
    for (uint64_t s = gw; s < p.n_snarls; s += stride, ++it) {
        const uint32_t stage = it % p.n_stages;
        const uint32_t parity = (it / p.n_stages) & 1u;
        // record of the snarl that will refill this stage: requested now, consumed after the compute
        const uint64_t s_next = s + uint64_t(p.n_stages) * stride;
        uint32_t next_word = (s_next < p.n_snarls) ? load_rec_word(p.rec, s_next, lane) : 0u;
        mbar_wait(&my_full[stage], parity);
        const uint8_t* blob = my_stages + size_t(stage) * stage_bytes;
        const uint8_t* rows_s = blob + kBlobBytes;
        const uint32_t hdr = *reinterpret_cast<const uint32_t*>(blob);
        const uint32_t np = hdr & 0xFFu;
        const uint32_t p0 = *reinterpret_cast<const uint32_t*>(blob + 4);
        if ((hdr >> 8 & 0xFFu) == 1u) {
            // ---- the common shape: two paths with two checked edges each (a biallelic snarl) ----------
            const uint32_t loc = uint32_t(blob[10]) | uint32_t(blob[11]) << 8 | uint32_t(blob[12]) << 16 |
                                 uint32_t(blob[13]) << 24;
            const uint32_t o00 = (loc & 0xFFu) * p.pitch, o01 = (loc >> 8 & 0xFFu) * p.pitch;
            const uint32_t o10 = (loc >> 16 & 0xFFu) * p.pitch, o11 = (loc >> 24) * p.pitch;
            uint32_t a0 = 0, c0 = 0, a1 = 0, c1 = 0;
            for (uint32_t c = lane; c < chunks; c += 32) {
                const uint4 m = *reinterpret_cast<const uint4*>(mask_s + c * 16);
                const uint4 v0 = and4(*reinterpret_cast<const uint4*>(rows_s + o00 + c * 16),
                                      *reinterpret_cast<const uint4*>(rows_s + o01 + c * 16));
                const uint4 v1 = and4(*reinterpret_cast<const uint4*>(rows_s + o10 + c * 16),
                                      *reinterpret_cast<const uint4*>(rows_s + o11 + c * 16));
                a0 += popc4(v0);
                c0 += popc4(and4(v0, m));
                a1 += popc4(v1);
                c1 += popc4(and4(v1, m));
            }
            if (packable) {
                const uint32_t r0 = __reduce_add_sync(0xffffffffu, a0 | c0 << 16);
                const uint32_t r1 = __reduce_add_sync(0xffffffffu, a1 | c1 << 16);
                a0 = r0 & 0xFFFFu; c0 = r0 >> 16;
                a1 = r1 & 0xFFFFu; c1 = r1 >> 16;
            } else {
                a0 = __reduce_add_sync(0xffffffffu, a0); c0 = __reduce_add_sync(0xffffffffu, c0);
                a1 = __reduce_add_sync(0xffffffffu, a1); c1 = __reduce_add_sync(0xffffffffu, c1);
            }
            if (lane < 4) {  // lane = path + 2 * (cases ? 1 : 0)
                const uint32_t a = (lane & 1u) ? a1 : a0, c = (lane & 1u) ? c1 : c0;
                uint32_t* dst = (lane & 2u) ? p.g1 : p.g0;
                dst[p0 + (lane & 1u)] = (lane & 2u) ? c : a - c;  // cases / controls (binary_table.cpp:35-39)
            }
        } else {
            // ---- any other staged shape: one path at a time ----------------------------------------------
            uint32_t off = kBlobHeader + np;
            for (uint32_t j = 0; j < np; ++j) {
                const uint32_t code = blob[kBlobHeader + j];
                uint32_t n_all = 0, n_case = 0;
                if (code == 0u) {  // no checked edge: every haplotype (snarl_analyzer.cpp:343-354)
                    n_all = p.n_hap;
                    n_case = p.n_case_hap;
                } else if (code != 0xFFu) {
                    if (code <= 4u) {
                        uint32_t roff[4];
#pragma unroll
                        for (int r = 0; r < 4; ++r) roff[r] = (uint32_t(r) < code) ? uint32_t(blob[off + r]) * p.pitch : 0u;
                        switch (code) {
                            case 1: path_counts<1>(rows_s, mask_s, roff, lane, chunks, n_all, n_case); break;
                            case 2: path_counts<2>(rows_s, mask_s, roff, lane, chunks, n_all, n_case); break;
                            case 3: path_counts<3>(rows_s, mask_s, roff, lane, chunks, n_all, n_case); break;
                            default: path_counts<4>(rows_s, mask_s, roff, lane, chunks, n_all, n_case); break;
                        }
                    } else {
                        for (uint32_t c = lane; c < chunks; c += 32) {
                            uint4 v = *reinterpret_cast<const uint4*>(rows_s + uint32_t(blob[off]) * p.pitch + c * 16);
                            for (uint32_t r = 1; r < code; ++r)
                                v = and4(v, *reinterpret_cast<const uint4*>(rows_s + uint32_t(blob[off + r]) * p.pitch + c * 16));
                            const uint4 m = *reinterpret_cast<const uint4*>(mask_s + c * 16);
                            n_all += popc4(v);
                            n_case += popc4(and4(v, m));
                        }
                    }
                    off += code;
                    if (packable) {
                        const uint32_t r = __reduce_add_sync(0xffffffffu, n_all | n_case << 16);
                        n_all = r & 0xFFFFu;
                        n_case = r >> 16;
                    } else {
                        n_all = __reduce_add_sync(0xffffffffu, n_all);
                        n_case = __reduce_add_sync(0xffffffffu, n_case);
                    }
                }
                if (lane == 0) {
                    p.g0[p0 + j] = n_all - n_case;  // controls  (binary_table.cpp:37-39)
                    p.g1[p0 + j] = n_case;          // cases     (binary_table.cpp:35-36)
                }
            }
        }
        // all lanes are done reading this stage: refill it with the snarl n_stages ahead
        __syncwarp();
        asm volatile("" : "+r"(next_word));  // keep the record's consumers below the compute
        if (s_next < p.n_snarls)
            issue_stage(next_word, lane, my_stages + size_t(stage) * stage_bytes, &my_full[stage], p.rows, p.blob,
                        p.pitch, policy);
    }
}

// =================================================================================================
// K1': straight from global memory (one warp per listed snarl)
// =================================================================================================
__global__ void __launch_bounds__(256) path_and_popc_ldg_kernel(const uint8_t* __restrict__ rows,
                                                                const uint8_t* __restrict__ case_mask, uint32_t pitch,
                                                                uint32_t n_hap, uint32_t n_case_hap, BatchDev b,
                                                                const uint32_t* __restrict__ list, uint64_t n_list,
                                                                uint32_t* __restrict__ g0, uint32_t* __restrict__ g1) {
    const uint32_t lane = threadIdx.x & 31;
    const uint64_t warp = (uint64_t(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const uint64_t n_warps = (uint64_t(gridDim.x) * blockDim.x) >> 5;
    const uint32_t chunks = pitch / 16;
    for (uint64_t i = warp; i < n_list; i += n_warps) {
        const uint64_t s = list ? list[i] : i;
        const uint64_t p0 = b.snarl_path_off[s], p1 = b.snarl_path_off[s + 1];
        for (uint64_t pth = p0; pth < p1; ++pth) {
            const uint64_t a = b.path_edge_off[pth], z = b.path_edge_off[pth + 1];
            bool missing = false;
            for (uint64_t e = a; e < z; ++e) missing |= (b.edge_row[e] == kMissing);
            uint32_t n_all = 0, n_case = 0;
            if (missing) {
            } else if (z == a) {
                n_all = n_hap;
                n_case = n_case_hap;
            } else {
                for (uint32_t c = lane; c < chunks; c += 32) {
                    uint4 v = ldg_stream(reinterpret_cast<const uint4*>(rows + size_t(b.edge_row[a]) * pitch) + c);
                    for (uint64_t e = a + 1; e < z; ++e)
                        v = and4(v, ldg_stream(reinterpret_cast<const uint4*>(rows + size_t(b.edge_row[e]) * pitch) + c));
                    const uint4 m = __ldg(reinterpret_cast<const uint4*>(case_mask) + c);
                    n_all += popc4(v);
                    n_case += popc4(and4(v, m));
                }
                n_all = __reduce_add_sync(0xffffffffu, n_all);
                n_case = __reduce_add_sync(0xffffffffu, n_case);
            }
            if (lane == 0) {
                g0[pth] = n_all - n_case;
                g1[pth] = n_case;
            }
        }
    }
}

// configuration of the TMA kernel for a given pitch / stage width; returns false if a single stage
// does not fit shared memory (then everything takes the LDG path)
struct K1Config {
    uint32_t n_warps, n_stages, stage_rows, smem_bytes;
};
static bool choose_k1_config(uint32_t pitch, uint32_t stage_rows, K1Config& cfg) {
    const uint32_t budget = 200 * 1024;
    if (stage_rows == 0) stage_rows = 1;
    const uint32_t stage_bytes = kBlobBytes + stage_rows * pitch;
    const uint32_t mask_bytes = (pitch + 127) & ~127u;
    if (mask_bytes + 2 * stage_bytes + 1024 > budget) return false;
    const uint32_t total_stages = (budget - mask_bytes - 1024) / stage_bytes;
    // measured on B200 (C3, profiles/): 13 warps x 2 stages 0.78 ms, 8 x 3 0.92 ms, 6 x 4 1.22 ms — a warp's compute
    // chain, not the copy depth, limits the rate, so shared memory goes to warps first
    uint32_t n_stages = 2;
    if (const char* e = std::getenv("STOAT_GPU_K1_STAGES")) {
        const int v = std::atoi(e);
        if (v >= 2 && v <= 8 && uint32_t(v) <= total_stages) n_stages = uint32_t(v);
    }
    uint32_t n_warps = total_stages / n_stages;
    if (const char* e = std::getenv("STOAT_GPU_K1_WARPS")) {
        const int v = std::atoi(e);
        if (v >= 1 && uint32_t(v) <= n_warps) n_warps = uint32_t(v);
    }
    if (n_warps > 16) {
        n_warps = 16;
        n_stages = total_stages / 16;
        if (n_stages > 4) n_stages = 4;
    }
    if (n_warps == 0) return false;
    cfg.n_warps = n_warps;
    cfg.n_stages = n_stages;
    cfg.stage_rows = stage_rows;
    const uint32_t bar_bytes = (n_warps * n_stages * 8 + 127) & ~127u;
    cfg.smem_bytes = bar_bytes + mask_bytes + n_warps * n_stages * stage_bytes;
    return true;
}

cudaError_t launch_path_and_popc(const MatrixDev& m, const uint8_t* case_mask, uint32_t n_case_hap,
                                     const BatchDev& b, uint32_t n_big, uint32_t max_stage_rows, uint32_t* counts,
                                     int num_sms, int variant, cudaStream_t st, uint32_t* launches, uint64_t s_begin,
                                     uint64_t s_count) {
    // [s_begin, s_begin + s_count): the snarls of this launch (the run is cut into chunks so that the statistics of a chunk
    // overlap the count pass of the next one, abi.cu); snarls on the LDG list are done with the first chunk
    if (b.n_snarls == 0 || s_count == 0) return cudaSuccess;
    if (s_begin >= b.n_snarls) return cudaSuccess;
    if (s_count > b.n_snarls - s_begin) s_count = b.n_snarls - s_begin;
    uint32_t* g0 = counts;
    uint32_t* g1 = counts + b.n_paths;
    const uint32_t pitch = uint32_t(m.pitch);
    K1Config cfg{};
    const bool use_tma = (variant == 0) && choose_k1_config(pitch, max_stage_rows, cfg) && (n_big < b.n_snarls);
    if (use_tma) {
        K1Params p{};
        p.rows = m.rows;
        p.case_mask = case_mask;
        p.rec = b.rec + s_begin;
        p.blob = b.blob;
        p.g0 = g0;
        p.g1 = g1;
        p.n_snarls = s_count;
        p.pitch = pitch;
        p.n_hap = uint32_t(m.n_cols);
        p.n_case_hap = n_case_hap;
        p.stage_rows = cfg.stage_rows;
        p.n_stages = cfg.n_stages;
        cudaError_t e = cudaFuncSetAttribute(path_and_popc_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             int(cfg.smem_bytes));
        if (e != cudaSuccess) return e;
        uint64_t want = (s_count + cfg.n_warps - 1) / cfg.n_warps;
        const unsigned grid = unsigned(want < uint64_t(num_sms) ? want : uint64_t(num_sms));
        path_and_popc_tma_kernel<<<grid, cfg.n_warps * 32, cfg.smem_bytes, st>>>(p);
        e = cudaGetLastError();
        if (e != cudaSuccess) return e;
        if (launches) ++*launches;
        if (n_big && s_begin == 0) {
            const unsigned blocks = unsigned(n_big < 148u * 8 ? (n_big + 7) / 8 : 148u * 8);
            path_and_popc_ldg_kernel<<<blocks, 256, 0, st>>>(m.rows, case_mask, pitch, uint32_t(m.n_cols), n_case_hap,
                                                             b, b.big_list, n_big, g0, g1);
            e = cudaGetLastError();
            if (e != cudaSuccess) return e;
            if (launches) ++*launches;
        }
        return cudaSuccess;
    }
    if (s_begin != 0) return cudaSuccess;  // LDG-only variant: everything went with the first chunk
    const uint64_t want = (b.n_snarls + 7) / 8;
    const unsigned blocks = unsigned(want < 148ull * 16 ? want : 148ull * 16);
    path_and_popc_ldg_kernel<<<blocks, 256, 0, st>>>(m.rows, case_mask, pitch, uint32_t(m.n_cols), n_case_hap, b,
                                                     nullptr, b.n_snarls, g0, g1);
    if (launches) ++*launches;
    return cudaGetLastError();
}

// =================================================================================================
// K2: per-snarl statistics
//   K2a binary_stats   one thread per snarl: column compaction, filter, chi2 (cheap, uniform work); 2x2 tables
//                      that need Fisher's test are appended to a work list
//   K2b fisher_exact   persistent lanes pull tables from that list: the recurrence runs 10^2..10^4 iterations
//                      depending on the table, so a static thread-per-snarl mapping leaves most lanes of a warp idle
//                      (measured: 8.3 of 32 lanes active); with dynamic refill every lane stays in the shared loop body
// =================================================================================================
__global__ void __launch_bounds__(128) binary_stats_kernel(BatchDev b, const uint32_t* __restrict__ g0v,
                                                           const uint32_t* __restrict__ g1v,
                                                           stoat_gpu_thresholds thr, uint64_t* __restrict__ total_sum,
                                                           uint8_t* __restrict__ filtered, double* __restrict__ p_chi2,
                                                           double* __restrict__ p_fisher,
                                                           uint8_t* __restrict__ chi2_code,
                                                           uint8_t* __restrict__ fisher_code,
                                                           uint4* __restrict__ fisher_tab,
                                                           uint32_t* __restrict__ fisher_idx,
                                                           uint32_t* __restrict__ fisher_cnt, uint64_t s_begin,
                                                           uint64_t s_end) {
    const uint64_t s = s_begin + uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    const uint32_t lane = threadIdx.x & 31;
    bool want_fisher = false;
    uint64_t c0[2] = {0, 0}, c1[2] = {0, 0};  // first two non-empty columns (the 2x2 case)
    if (s < s_end) {
        const uint64_t p0 = b.snarl_path_off[s], p1 = b.snarl_path_off[s + 1];
        // pass 1 over the columns: totals, compaction count, MAF count (snarl_analyzer.cpp:686-732)
        uint64_t tot = 0, row0 = 0, row1 = 0;
        uint32_t ncol = 0, above = 0;
        bool col_low32_zero = false;  // chi2_2xN's any_of takes the column total as int (:233)
        for (uint64_t p = p0; p < p1; ++p) {
            const uint64_t a = g0v[p], c = g1v[p];
            const uint64_t cs = a + c;
            if (cs == 0) continue;  // remove_empty_columns_binary_table
            if (ncol < 2) {
                c0[ncol] = a;
                c1[ncol] = c;
            }
            ++ncol;
            tot += cs;
            row0 += a;
            row1 += c;
            if (uint32_t(cs) == 0) col_low32_zero = true;
            const double freq1 = __ddiv_rn(double(c), double(cs));
            const double maf = fmin(freq1, __dsub_rn(1.0, freq1));
            if (maf > thr.maf_threshold) ++above;
        }
        total_sum[s] = tot;
        // filtration_binary_table (:714, :731)
        const bool filt = (tot / 2 < thr.min_individuals) || (tot < thr.min_haplotypes) || (ncol < 2) || (above < 2);
        filtered[s] = filt ? 1 : 0;
        PVal chi{0.0, kPNa};
        if (!filt) {
            if (ncol == 2) {  // fisher_khi2 (:369-375): a = g0[0], b = g0[1], c = g1[0], d = g1[1]
                chi = chi2_2x2(c0[0], c0[1], c1[0], c1[1]);
                // fastFishersExactTest's own NA rule (:268-270); everything else goes to the work list
                const uint64_t m11 = c0[0], m12 = c0[1], m21 = c1[0], m22 = c1[1];
                want_fisher = !((m11 | m12) == 0 || (m21 | m22) == 0 || (m11 | m21) == 0 || (m12 | m22) == 0);
            } else {  // chi2_2xN (:214-257)
                if (!(tot == 0 || row0 == 0 || row1 == 0 || col_low32_zero)) {
                    double stat = 0.0;
                    for (uint64_t p = p0; p < p1; ++p) {
                        const uint64_t a = g0v[p], c = g1v[p];
                        if (a + c == 0) continue;
                        stat = chi2_2xN_column(stat, a, c, row0, row1, tot);
                    }
                    chi = {chi2_pvalue_like_reference(stat, ncol - 1), kPValue};
                }
            }
        }
        p_chi2[s] = chi.v;
        chi2_code[s] = chi.code;
        p_fisher[s] = 0.0;
        fisher_code[s] = kPNa;
    }
    // warp-aggregated append to the Fisher work list
    const uint32_t need = __ballot_sync(0xffffffffu, want_fisher);
    if (need) {
        const int leader = __ffs(need) - 1;
        uint32_t base = 0;
        if (int(lane) == leader) base = atomicAdd(fisher_cnt, uint32_t(__popc(need)));
        base = __shfl_sync(0xffffffffu, base, leader);
        if (want_fisher) {
            const uint32_t slot = base + __popc(need & ((1u << lane) - 1u));
            fisher_tab[slot] = make_uint4(uint32_t(c0[0]), uint32_t(c0[1]), uint32_t(c1[0]), uint32_t(c1[1]));
            fisher_idx[slot] = uint32_t(s);
        }
    }
}

// K2b.  One lane = one running recurrence (the same arithmetic as fisher_2x2 in device_math.cuh, operation for
// operation, split at the loop body so that lanes can be refilled from the work list).  The body of all three loops of
// fastFishersExactTest (src/stats_test.cpp:306-357) is
//     d1 += 1; d2 += 1; cur_prob *= (n1 * n2) / (d1 * d2); n1 -= 1; n2 -= 1;
// The products are carried instead of the factors:  N = n1 * n2 and D = d1 * d2 are integers below 2^53 (every count
// is below 2^26 here; larger tables take the factor form), so  D' = D + (d1 + d2 + 1)  and  N' = N - (n1 + n2 - 1)
// are the exactly representable values the reference's two multiplications round to — bit-identical, two DMUL fewer.
// The running product and the tail sums keep the reference's order of operations (no FMA contraction); the quotient
// N / D is N * fast_rcp(D) (below): within ~1e-13 of the reference over a table, inside its own tie margin.

// 1 / d for an exact positive integer d < 2^53 to ~1 ulp: MUFU.RCP64H seed + two Newton steps (5 instructions; the
// correctly rounded __ddiv_rn is ~3x that plus a branch).  The recurrence then differs from the reference's
// `cur_prob *= (n1 * n2) / (d1 * d2)` by ~1e-16 per step, ~1e-13 over a table — four orders below the 1e-9 budget, and
// below the reference's own tie margin kExactTestEpsilon2 = 9.1e-13 (stats_test.hpp:59), so tables tied with the observed
// one fall on the same side of `cur_prob < kExactTestBias` as in the reference.
__device__ __forceinline__ double fast_rcp(double d) {
    double x;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(x) : "d"(d));
    double e = fma(-d, x, 1.0);
    x = fma(x, e, x);
    e = fma(-d, x, 1.0);
    return fma(x, e, x);
}

template <bool kFastRcp>
__global__ void __launch_bounds__(128) fisher_exact_kernel(const uint4* __restrict__ tab,
                                                           const uint32_t* __restrict__ idx,
                                                           uint32_t* __restrict__ counters,  // [0] n_work, [1] next
                                                           double* __restrict__ p_fisher,
                                                           uint8_t* __restrict__ fisher_code) {
    constexpr double kEps2 = 9.094947017729282e-13;
    constexpr double kBias = 1.0339757656912846e-25;
    constexpr double kStart = (1 - kEps2) * kBias;
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t n_work = counters[0];
    bool active = false, drained = false;
    uint32_t snarl = 0, a11 = 0, a12 = 0, a21 = 0, a22 = 0, rem = 0;
    int phase = 0;
    double N = 0, sN = 0, D = 0, sD = 0, cur = 0, tprob = 0, cprob = 0;
    auto finish = [&](double v, uint8_t code) {
        p_fisher[snarl] = v;
        fisher_code[snarl] = code;
        active = false;
    };
    auto start_near_tail = [&]() {  // :339-344, or the final return when m11 == 0 (:359)
        if (a11 == 0) {
            finish(__ddiv_rn(tprob, __dadd_rn(cprob, tprob)), kPValue);
            return;
        }
        N = __dmul_rn(double(a11), double(a22));
        sN = double(a11) + double(a22) - 1.0;
        D = __dmul_rn(double(a12), double(a21));
        sD = double(a12) + double(a21) + 1.0;
        cur = kStart;
        rem = a11;
        phase = 3;
    };
    for (;;) {
        const uint32_t need = __ballot_sync(0xffffffffu, !active && !drained);
        if (need) {
            const int leader = __ffs(need) - 1;
            uint32_t base = 0;
            if (int(lane) == leader) base = atomicAdd(&counters[1], uint32_t(__popc(need)));
            base = __shfl_sync(0xffffffffu, base, leader);
            if (!active && !drained) {
                const uint32_t w = base + __popc(need & ((1u << lane) - 1u));
                if (w < n_work) {
                    const uint4 t = tab[w];
                    snarl = idx[w];
                    if ((t.x | t.y | t.z | t.w) >> 26) {  // products would leave the exact-integer range
                        const PVal r = fisher_2x2(t.x, t.y, t.z, t.w);
                        p_fisher[snarl] = r.v;
                        fisher_code[snarl] = r.code;
                    } else {
                        a11 = t.x; a12 = t.y; a21 = t.z; a22 = t.w;
                        uint32_t u;
                        if (a12 > a21) { u = a12; a12 = a21; a21 = u; }  // stats_test.cpp:280-284
                        if (a11 > a22) { u = a11; a11 = a22; a22 = u; }  // :286-290
                        if (uint64_t(a11) * a22 > uint64_t(a12) * a21) {  // :292-299
                            u = a11; a11 = a12; a12 = u;
                            u = a21; a21 = a22; a22 = u;
                        }
                        N = __dmul_rn(double(a12), double(a21));
                        sN = double(a12) + double(a21) - 1.0;
                        D = __dmul_rn(double(a11), double(a22));
                        sD = double(a11) + double(a22) + 1.0;
                        tprob = kStart; cur = kStart; cprob = 0.0;
                        rem = a12;
                        phase = 1;
                        active = true;
                    }
                } else {
                    drained = true;
                }
            }
        }
        if (!__any_sync(0xffffffffu, active)) break;
        {
            if (active && rem == 0) {  // the loop of the current phase ran out before a body execution (:306, :326)
                if (phase == 1 && cprob == 0) finish(1.0, kPOne);  // :322-324
                else start_near_tail();
            }
            if (active) {
                D = __dadd_rn(D, sD);
                sD += 2.0;
                cur = kFastRcp ? __dmul_rn(cur, __dmul_rn(N, fast_rcp(D))) : __dmul_rn(cur, __ddiv_rn(N, D));
                N = __dsub_rn(N, sN);
                sN -= 2.0;
                --rem;
                if (phase == 1) {
                    if (cur > DBL_MAX) {  // :312-314
                        finish(0.0, kPZero);
                    } else if (cur < kBias) {  // :315-318
                        tprob = __dadd_rn(tprob, cur);
                        if (cprob == 0) finish(1.0, kPOne);
                        else phase = 2;
                    } else {
                        cprob = __dadd_rn(cprob, cur);
                    }
                } else {
                    const double preaddp = tprob;
                    tprob = __dadd_rn(tprob, cur);
                    if (tprob <= preaddp) {
                        if (phase == 2) start_near_tail();  // :334-336
                        else finish(__ddiv_rn(preaddp, __dadd_rn(cprob, preaddp)), kPValue);  // :352-354
                    } else if (phase == 3 && rem == 0) {
                        finish(__ddiv_rn(tprob, __dadd_rn(cprob, tprob)), kPValue);  // :356, :359
                    }
                }
            }
        }
    }
}

// statistics of the snarls [s_begin, s_begin + s_count); the Fisher work list of the range lives at fisher_tab / fisher_idx
// + s_begin (at most one table per snarl) with its own pair of counters
cudaError_t launch_binary_stats(const BatchDev& b, const uint32_t* counts, const stoat_gpu_thresholds& thr,
                                uint64_t* total_sum, uint8_t* filtered, double* p_chi2, double* p_fisher,
                                uint8_t* chi2_code, uint8_t* fisher_code, uint4* fisher_tab, uint32_t* fisher_idx,
                                uint32_t* fisher_counters, int num_sms, cudaStream_t st, uint64_t s_begin,
                                uint64_t s_count) {
    if (b.n_snarls == 0 || s_count == 0) return cudaSuccess;
    if (s_begin >= b.n_snarls) return cudaSuccess;
    if (s_count > b.n_snarls - s_begin) s_count = b.n_snarls - s_begin;
    cudaError_t e = cudaMemsetAsync(fisher_counters, 0, 8, st);
    if (e != cudaSuccess) return e;
    const unsigned blocks = unsigned((s_count + 127) / 128);
    binary_stats_kernel<<<blocks, 128, 0, st>>>(b, counts, counts + b.n_paths, thr, total_sum, filtered, p_chi2,
                                                p_fisher, chi2_code, fisher_code, fisher_tab + s_begin,
                                                fisher_idx + s_begin, fisher_counters, s_begin, s_begin + s_count);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    const uint64_t want = (s_count + 127) / 128;
    static const int per_sm = [] {
        const char* v = std::getenv("STOAT_GPU_FISHER_BLOCKS");
        const int n = v ? std::atoi(v) : 0;
        return n >= 1 ? n : 8;
    }();
    const unsigned fblocks = unsigned(want < uint64_t(num_sms) * per_sm ? want : uint64_t(num_sms) * per_sm);
    // STOAT_GPU_FISHER=exact: the quotient by __ddiv_rn, bit-identical to the reference's recurrence (15 % slower)
    const char* fv = std::getenv("STOAT_GPU_FISHER");
    if (fv && fv[0] == 'e')
        fisher_exact_kernel<false><<<fblocks, 128, 0, st>>>(fisher_tab + s_begin, fisher_idx + s_begin, fisher_counters,
                                                            p_fisher, fisher_code);
    else
        fisher_exact_kernel<true><<<fblocks, 128, 0, st>>>(fisher_tab + s_begin, fisher_idx + s_begin, fisher_counters,
                                                           p_fisher, fisher_code);
    return cudaGetLastError();
}

}  // namespace stoat
