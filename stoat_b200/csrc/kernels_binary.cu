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
    uint32_t chunk_bytes;  // bytes of every row one stage holds (== pitch when a stage holds whole rows)
    uint32_t n_chunks;     // column chunks per row: ceil(pitch / chunk_bytes)
};

// AND of R staged rows for one 16-byte vector, then the two popcounts.  R is a compile-time constant so the loads and
// ANDs are straight-line code.  `nvec` 16-byte vectors of this column chunk, `mask_c` the case mask at the chunk's offset.
template <int R>
__device__ __forceinline__ void path_counts(const uint8_t* rows_s, const uint8_t* mask_c, const uint32_t (&roff)[4],
                                            uint32_t lane, uint32_t nvec, uint32_t& n_all, uint32_t& n_case) {
    for (uint32_t c = lane; c < nvec; c += 32) {
        uint4 v = *reinterpret_cast<const uint4*>(rows_s + roff[0] + c * 16);
#pragma unroll
        for (int r = 1; r < R; ++r) v = and4(v, *reinterpret_cast<const uint4*>(rows_s + roff[r] + c * 16));
        const uint4 m = *reinterpret_cast<const uint4*>(mask_c + c * 16);
        n_all += popc4(v);
        n_case += popc4(and4(v, m));
    }
}

// One warp = one pipeline.  A work item is (snarl, column chunk): the stage receives the snarl's path program and the
// bytes [chunk * chunk_bytes, +chunk_bytes) of each of its distinct rows, so the shared memory a stage needs — and with it
// the number of warps and copies in flight per SM — does not depend on the haplotype count (C3: 1,280-byte rows, one
// chunk; C5: 6,272-byte rows, four chunks of 1,568 bytes).  A warp walks the chunks of a snarl in order and carries the
// per-lane popcounts across them; kMulti = false is the one-chunk case with the hand-over code compiled out.
template <bool kMulti>
__device__ __forceinline__ void path_and_popc_body(const K1Params& p) {
    extern __shared__ __align__(128) uint8_t smem[];
    const uint32_t n_warps = blockDim.x >> 5;
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t slot = kMulti ? p.chunk_bytes : p.pitch;  // bytes of a row in a stage
    const uint32_t n_chunks = kMulti ? p.n_chunks : 1u;
    const uint32_t stage_bytes = kBlobBytes + p.stage_rows * slot;
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
    const uint64_t policy = make_evict_first_policy();
    const bool packable = p.n_hap < 65536u;  // both counts of a path fit one 32-bit warp reduction

    // issue cursor: the (snarl, chunk) the next refill loads
    uint64_t s_iss = gw;
    uint32_t c_iss = 0;
    auto issue = [&](uint32_t rec_word, uint32_t stage) {
        if (kMulti) {
            const uint32_t off = c_iss * slot;
            const uint32_t bytes = (p.pitch - off < slot) ? p.pitch - off : slot;
            issue_stage_slice(rec_word, lane, my_stages + size_t(stage) * stage_bytes, &my_full[stage], p.rows, p.blob,
                              p.pitch, off, bytes, slot, policy);
            if (++c_iss == n_chunks) {
                c_iss = 0;
                s_iss += stride;
            }
        } else {
            issue_stage(rec_word, lane, my_stages + size_t(stage) * stage_bytes, &my_full[stage], p.rows, p.blob, p.pitch,
                        policy);
            s_iss += stride;
        }
    };

    // prologue: fill the pipeline
    for (uint32_t st = 0; st < p.n_stages && s_iss < p.n_snarls; ++st) issue(load_rec_word(p.rec, s_iss, lane), st);

    uint32_t it = 0, chunk = 0;
    uint32_t a0 = 0, c0 = 0, a1 = 0, c1 = 0;  // the biallelic shape's per-lane counts, carried over a snarl's chunks
    for (uint64_t s = gw; s < p.n_snarls; ++it) {
        const uint32_t stage = it % p.n_stages;
        const uint32_t parity = (it / p.n_stages) & 1u;
        // record of the snarl that will refill this stage: requested now, consumed after the compute
        const bool more = s_iss < p.n_snarls;
        uint32_t next_word = more ? load_rec_word(p.rec, s_iss, lane) : 0u;
        mbar_wait(&my_full[stage], parity);
        const uint8_t* blob = my_stages + size_t(stage) * stage_bytes;
        const uint8_t* rows_s = blob + kBlobBytes;
        const uint32_t hdr = *reinterpret_cast<const uint32_t*>(blob);
        const uint32_t np = hdr & 0xFFu;
        const uint32_t p0 = *reinterpret_cast<const uint32_t*>(blob + 4);
        const uint32_t col0 = chunk * slot;  // byte offset of this chunk in a row / in the case mask
        const uint32_t nvec = ((p.pitch - col0 < slot) ? p.pitch - col0 : slot) / 16;
        const uint8_t* mask_c = mask_s + col0;
        const bool first = !kMulti || chunk == 0, last = !kMulti || chunk + 1 == n_chunks;
        if ((hdr >> 8 & 0xFFu) == 1u) {
            // ---- the common shape: two paths with two checked edges each (a biallelic snarl) ----------
            const uint32_t loc = uint32_t(blob[10]) | uint32_t(blob[11]) << 8 | uint32_t(blob[12]) << 16 |
                                 uint32_t(blob[13]) << 24;
            const uint32_t o00 = (loc & 0xFFu) * slot, o01 = (loc >> 8 & 0xFFu) * slot;
            const uint32_t o10 = (loc >> 16 & 0xFFu) * slot, o11 = (loc >> 24) * slot;
            if (first) a0 = c0 = a1 = c1 = 0;
            for (uint32_t c = lane; c < nvec; c += 32) {
                const uint4 m = *reinterpret_cast<const uint4*>(mask_c + c * 16);
                const uint4 v0 = and4(*reinterpret_cast<const uint4*>(rows_s + o00 + c * 16),
                                      *reinterpret_cast<const uint4*>(rows_s + o01 + c * 16));
                const uint4 v1 = and4(*reinterpret_cast<const uint4*>(rows_s + o10 + c * 16),
                                      *reinterpret_cast<const uint4*>(rows_s + o11 + c * 16));
                a0 += popc4(v0);
                c0 += popc4(and4(v0, m));
                a1 += popc4(v1);
                c1 += popc4(and4(v1, m));
            }
            if (last) {
                uint32_t ra0, rc0, ra1, rc1;
                if (packable) {
                    const uint32_t r0 = __reduce_add_sync(0xffffffffu, a0 | c0 << 16);
                    const uint32_t r1 = __reduce_add_sync(0xffffffffu, a1 | c1 << 16);
                    ra0 = r0 & 0xFFFFu; rc0 = r0 >> 16;
                    ra1 = r1 & 0xFFFFu; rc1 = r1 >> 16;
                } else {
                    ra0 = __reduce_add_sync(0xffffffffu, a0); rc0 = __reduce_add_sync(0xffffffffu, c0);
                    ra1 = __reduce_add_sync(0xffffffffu, a1); rc1 = __reduce_add_sync(0xffffffffu, c1);
                }
                if (lane < 4) {  // lane = path + 2 * (cases ? 1 : 0)
                    const uint32_t a = (lane & 1u) ? ra1 : ra0, c = (lane & 1u) ? rc1 : rc0;
                    uint32_t* dst = (lane & 2u) ? p.g1 : p.g0;
                    dst[p0 + (lane & 1u)] = (lane & 2u) ? c : a - c;  // cases / controls (binary_table.cpp:35-39)
                }
            }
        } else {
            // ---- any other staged shape: one path at a time; with several chunks lane 0 adds each chunk's counts to
            // what the earlier chunks of the snarl left in g0 / g1 (same thread, program order) ----------------------
            uint32_t off = kBlobHeader + np;
            for (uint32_t j = 0; j < np; ++j) {
                const uint32_t code = blob[kBlobHeader + j];
                uint32_t n_all = 0, n_case = 0;
                if (code == 0u) {  // no checked edge: every haplotype (snarl_analyzer.cpp:343-354)
                    if (first) {
                        n_all = p.n_hap;
                        n_case = p.n_case_hap;
                    }
                } else if (code != 0xFFu) {
                    if (code <= 4u) {
                        uint32_t roff[4];
#pragma unroll
                        for (int r = 0; r < 4; ++r) roff[r] = (uint32_t(r) < code) ? uint32_t(blob[off + r]) * slot : 0u;
                        switch (code) {
                            case 1: path_counts<1>(rows_s, mask_c, roff, lane, nvec, n_all, n_case); break;
                            case 2: path_counts<2>(rows_s, mask_c, roff, lane, nvec, n_all, n_case); break;
                            case 3: path_counts<3>(rows_s, mask_c, roff, lane, nvec, n_all, n_case); break;
                            default: path_counts<4>(rows_s, mask_c, roff, lane, nvec, n_all, n_case); break;
                        }
                    } else {
                        for (uint32_t c = lane; c < nvec; c += 32) {
                            uint4 v = *reinterpret_cast<const uint4*>(rows_s + uint32_t(blob[off]) * slot + c * 16);
                            for (uint32_t r = 1; r < code; ++r)
                                v = and4(v, *reinterpret_cast<const uint4*>(rows_s + uint32_t(blob[off + r]) * slot + c * 16));
                            const uint4 m = *reinterpret_cast<const uint4*>(mask_c + c * 16);
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
                    uint32_t ctl = n_all - n_case, cas = n_case;  // controls / cases (binary_table.cpp:35-39)
                    if (!first) {
                        ctl += p.g0[p0 + j];
                        cas += p.g1[p0 + j];
                    }
                    p.g0[p0 + j] = ctl;
                    p.g1[p0 + j] = cas;
                }
            }
        }
        // all lanes are done reading this stage: refill it with the item n_stages ahead
        __syncwarp();
        asm volatile("" : "+r"(next_word));  // keep the record's consumers below the compute
        if (more) issue(next_word, stage);
        if (kMulti) {
            if (++chunk == n_chunks) {
                chunk = 0;
                s += stride;
            }
        } else {
            s += stride;
        }
    }
}

__global__ void __launch_bounds__(512, 1) path_and_popc_tma_kernel(const K1Params p) { path_and_popc_body<false>(p); }
__global__ void __launch_bounds__(512, 1) path_and_popc_tma_chunked_kernel(const K1Params p) { path_and_popc_body<true>(p); }
// The same two kernels held to 64 registers per thread: 416 threads x 64 leave room in an SM's register file for four
// fisher_exact blocks next to the persistent count pass (the overlapped run of abi.cu: FP64 work under the HBM stream).
__global__ void __maxnreg__(64) path_and_popc_tma_lean_kernel(const K1Params p) {
    path_and_popc_body<false>(p);
}
__global__ void __maxnreg__(64) path_and_popc_tma_chunked_lean_kernel(const K1Params p) {
    path_and_popc_body<true>(p);
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
    uint32_t n_warps, n_stages, stage_rows, smem_bytes, chunk_bytes, n_chunks;
};
static int k1_env(const char* name, int dflt) {
    const char* e = std::getenv(name);
    return e ? std::atoi(e) : dflt;
}
static bool choose_k1_config(uint32_t pitch, uint32_t stage_rows, K1Config& cfg) {
    const uint32_t budget = 200 * 1024;
    if (stage_rows == 0) stage_rows = 1;
    // Rows wider than 2 KB are staged in column chunks of about 1.5 KB, so that the stage stays the size it has at C3
    // (1,280-byte rows) and the SM keeps ~10-13 warps x 2 stages in flight whatever the haplotype count.  Whole-row stages
    // at H = 50,000 (6,272-byte rows) would leave 2 warps per SM.
    uint32_t target = uint32_t(k1_env("STOAT_GPU_K1_CHUNK", 1536));
    if (target < 256) target = 256;
    target = (target + 15u) & ~15u;
    uint32_t n_chunks = 1, chunk_bytes = pitch;
    if (pitch > 2048 || pitch > target + target / 3) {
        n_chunks = (pitch + target - 1) / target;
        chunk_bytes = (((pitch + n_chunks - 1) / n_chunks) + 15u) & ~15u;
        n_chunks = (pitch + chunk_bytes - 1) / chunk_bytes;
    }
    const uint32_t stage_bytes = kBlobBytes + stage_rows * chunk_bytes;
    const uint32_t mask_bytes = (pitch + 127) & ~127u;
    if (mask_bytes + 2 * stage_bytes + 1024 > budget) return false;
    const uint32_t total_stages = (budget - mask_bytes - 1024) / stage_bytes;
    // measured on B200 (C3, profiles/): 13 warps x 2 stages 0.78 ms, 8 x 3 0.92 ms, 6 x 4 1.22 ms — a warp's compute
    // chain, not the copy depth, limits the rate, so shared memory goes to warps first
    uint32_t n_stages = 2;
    {
        const int v = k1_env("STOAT_GPU_K1_STAGES", 0);
        if (v >= 2 && v <= 8 && uint32_t(v) <= total_stages) n_stages = uint32_t(v);
    }
    uint32_t n_warps = total_stages / n_stages;
    {
        const int v = k1_env("STOAT_GPU_K1_WARPS", 0);
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
    cfg.chunk_bytes = chunk_bytes;
    cfg.n_chunks = n_chunks;
    const uint32_t bar_bytes = (n_warps * n_stages * 8 + 127) & ~127u;
    cfg.smem_bytes = bar_bytes + mask_bytes + n_warps * n_stages * stage_bytes;
    return true;
}

cudaError_t launch_path_and_popc(const MatrixDev& m, const uint8_t* case_mask, uint32_t n_case_hap,
                                     const BatchDev& b, uint32_t n_big, uint32_t max_stage_rows, uint32_t* counts,
                                     int num_sms, int variant, cudaStream_t st, uint32_t* launches, uint64_t s_begin,
                                     uint64_t s_count, bool lean_registers) {
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
        p.chunk_bytes = cfg.chunk_bytes;
        p.n_chunks = cfg.n_chunks;
        using K1Kernel = void (*)(const K1Params);
        const bool multi = cfg.n_chunks > 1;
        K1Kernel fn = lean_registers ? (multi ? path_and_popc_tma_chunked_lean_kernel : path_and_popc_tma_lean_kernel)
                                     : (multi ? path_and_popc_tma_chunked_kernel : path_and_popc_tma_kernel);
        cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, int(cfg.smem_bytes));
        if (e != cudaSuccess) return e;
        uint64_t want = (s_count + cfg.n_warps - 1) / cfg.n_warps;
        const unsigned grid = unsigned(want < uint64_t(num_sms) ? want : uint64_t(num_sms));
        fn<<<grid, cfg.n_warps * 32, cfg.smem_bytes, st>>>(p);
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

// Control flow.  A table passes through three loops (phase 1: towards the centre until the term drops below the bias,
// 2: the rest of that tail, 3: the near tail from the observed table outwards), ~380 iterations in all at H = 10k, and
// has exactly two or three "events": the hand-over 2 -> 3, the end, and the rare early exits.  The loop below keeps every
// lane in ONE branch-free body — the phase only selects which running sum the term goes to, by predication — and leaves
// it when some lane reports an event; events and refills from the work list are handled in the (divergent) code behind
// it.  The first version ran each phase's bookkeeping as its own branch and polled the work list in every iteration:
// 21 of 32 lanes active, more than half of the issued instructions control flow (profiles/r1g_fisher.txt).
// Positive doubles order like their bit patterns, so the two comparisons of the body are integer compares and stay off
// the FP64 pipe, which is what bounds this kernel.
// tables with a count of 2^26 or more: the factor form of device_math.cuh, kept out of line (never taken below 6.7e7 haplotypes)
__device__ __noinline__ void fisher_big_table(uint4 t, uint32_t snarl, double* p_fisher, uint8_t* fisher_code) {
    const PVal r = fisher_2x2(t.x, t.y, t.z, t.w);
    p_fisher[snarl] = r.v;
    fisher_code[snarl] = r.code;
}

template <bool kFastRcp>
__global__ void __launch_bounds__(128, 8) fisher_exact_kernel(const uint4* __restrict__ tab,
                                                           const uint32_t* __restrict__ idx,
                                                           uint32_t* __restrict__ counters,  // [0] n_work, [1] next
                                                           double* __restrict__ p_fisher,
                                                           uint8_t* __restrict__ fisher_code) {
    constexpr double kEps2 = 9.094947017729282e-13;
    constexpr double kBias = 1.0339757656912846e-25;
    constexpr double kStart = (1 - kEps2) * kBias;
    const long long kBiasBits = __double_as_longlong(kBias);
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t n_work = counters[0];
    bool active = false, drained = false;
    uint32_t snarl = 0, a11 = 0, a12 = 0, a21 = 0, a22 = 0, rem = 1;
    int phase = 2;
    // an idle lane runs the body on these values: finite, no event
    double N = 0, sN = 0, D = 1, sD = 0, cur = 0, tprob = 1, cprob = 0;
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
        // ---- refill the idle lanes from the work list ----------------------------------------------------------
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
                        fisher_big_table(t, snarl, p_fisher, fisher_code);
                    } else {
                        a11 = t.x; a12 = t.y; a21 = t.z; a22 = t.w;
                        uint32_t u;
                        if (a12 > a21) { u = a12; a12 = a21; a21 = u; }  // stats_test.cpp:280-284
                        if (a11 > a22) { u = a11; a11 = a22; a22 = u; }  // :286-290
                        if (uint64_t(a11) * a22 > uint64_t(a12) * a21) {  // :292-299
                            u = a11; a11 = a12; a12 = u;
                            u = a21; a21 = a22; a22 = u;
                        }
                        if (a12 == 0) {  // the first loop does not run (:306), cprob == 0 (:322-324)
                            p_fisher[snarl] = 1.0;
                            fisher_code[snarl] = kPOne;
                        } else {
                            N = __dmul_rn(double(a12), double(a21));
                            sN = double(a12) + double(a21) - 1.0;
                            D = __dmul_rn(double(a11), double(a22));
                            sD = double(a11) + double(a22) + 1.0;
                            tprob = kStart; cur = kStart; cprob = 0.0;
                            rem = a12;
                            phase = 1;
                            active = true;
                        }
                    }
                } else {
                    drained = true;
                }
            }
            if (__any_sync(0xffffffffu, !active && !drained)) continue;  // a lane drew a table that was done at once
        }
        if (!__any_sync(0xffffffffu, active)) break;
        if (!active) {  // parked: a lane whose list ran dry iterates on harmless values
            N = 0; sN = 0; D = 1; sD = 0; cur = 0; tprob = 1; cprob = 0; rem = 0x7FFFFFFFu; phase = 2;
        }
        // ---- the recurrence: every lane in the same straight-line body until some lane has an event ---------------
        bool ev, to_t;
        double pre;
        do {
            D = __dadd_rn(D, sD);
            sD += 2.0;
            cur = kFastRcp ? __dmul_rn(cur, __dmul_rn(N, fast_rcp(D))) : __dmul_rn(cur, __ddiv_rn(N, D));
            N = __dsub_rn(N, sN);
            sN -= 2.0;
            --rem;
            const bool p1 = phase == 1;
            // cur >= +0: `cur < kBias` (:315) on the bit patterns; +inf and NaN patterns are above kBias's
            to_t = !p1 || __double_as_longlong(cur) < kBiasBits;
            pre = tprob;
            if (to_t) tprob = __dadd_rn(tprob, cur);   // :316 / :331-332 / :349-350
            else cprob = __dadd_rn(cprob, cur);        // :320
            // phase 1: overflow (:312), the term fell below the bias (:315), or the loop ran out;
            // phases 2, 3: the sum stopped growing (:333 / :351: tprob <= preaddp, i.e. == for a term >= 0), or ran out
            const bool stalled = __double_as_longlong(tprob) == __double_as_longlong(pre);
            ev = (p1 ? to_t || __double2hiint(cur) >= 0x7FF00000 : stalled) || rem == 0;
            ev = ev && active;
        } while (!__any_sync(0xffffffffu, ev));
        // ---- events --------------------------------------------------------------------------------------------------
        if (ev) {
            if (phase == 1) {
                if (cur > DBL_MAX) finish(0.0, kPZero);            // :312-314
                else if (cprob == 0) finish(1.0, kPOne);           // :322-324 (only after a break: cprob grows otherwise)
                else if (to_t && rem != 0) phase = 2;              // :315-318, on to the second loop
                else start_near_tail();                            // the first loop used up the column: the second is empty
            } else if (phase == 2) {
                start_near_tail();                                 // :334-336 or ran out
            } else if (__double_as_longlong(tprob) == __double_as_longlong(pre)) {
                finish(__ddiv_rn(pre, __dadd_rn(cprob, pre)), kPValue);    // :352-354
            } else {
                finish(__ddiv_rn(tprob, __dadd_rn(cprob, tprob)), kPValue);  // :356, :359
            }
        }
    }
}

// K2b, block form (default).  The serial kernel above issues ~34 instructions per step of the recurrence, 13 of them on the
// FP64 pipe (the quotient alone is MUFU + 4 DFMA + DMUL), and a table has three events that run divergent code.  Here a lane
// advances its table M steps at a time without a division and without a per-step decision:
//     N_j, D_j   the carried integer products of the serial form (exact)
//     PN_j = (cur / bias) * N_1 ... N_j,   PD_j = D_1 ... D_j            =>  term_j = bias * PN_j / PD_j
//     A_j  = A_{j-1} * D_j + PN_j                                         =>  sum of the block's terms = bias * A_M / PD_M
// (Horner over the common denominator: M steps cost 4 DADD + 2 DMUL + 1 DFMA each and ONE reciprocal.)  Which running sum
// a term belongs to is the reference's own test `cur_prob < kExactTestBias` (src/stats_test.cpp:315), here PN_j < PD_j on
// the bit patterns; the terms below the bias are summed a second time (T_j), so the block adds bias * (A - T) / PD to cprob
// and bias * T / PD to tprob.  The sequence of terms is unimodal, so that test is also what separates the reference's first
// loop (towards and past the centre, :306-321) from its second (:326-337): the two become ONE pass over the far side with no
// hand-over event, followed by the near side (:339-357).  A side ends when its last term no longer changes tprob
// (:333, :351: here term < tprob * 2^-53 at the end of a block — at most M - 1 terms, each below an ulp of tprob, are
// added beyond the reference's stopping point) or when the column is used up (:306 `cur12 > 0.5`); steps past the end of
// the column multiply by N = 0 and add exact zeros.  A block whose products leave the double range
// sends the table through the serial routine of device_math.cuh.  Not bit-identical to the reference (sums are associated
// differently, ~1e-15 relative); STOAT_GPU_FISHER=exact is.  Ties within kExactTestEpsilon2 = 9.1e-13 of the observed
// table's probability fall on the reference's side: the block form's error is three orders below that margin.
__device__ unsigned long long g_fisher_dbg[4];  // STOAT_GPU_FISHER_DEBUG: fallbacks, lane-blocks, tables, far-side blocks
template <int M>
__global__ void __launch_bounds__(128, 8) fisher_block_kernel(const uint4* __restrict__ tab,
                                                              const uint32_t* __restrict__ idx,
                                                              uint32_t* __restrict__ counters,  // [0] n_work, [1] next
                                                              double* __restrict__ p_fisher,
                                                              uint8_t* __restrict__ fisher_code) {
    constexpr double kEps2 = 9.094947017729282e-13;
    constexpr double kBias = 1.0339757656912846e-25;
    constexpr double kStart = (1 - kEps2) * kBias;
    constexpr double kInvBias = 1.0 / kBias;
    const long long kBiasBits = __double_as_longlong(kBias);
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t n_work = counters[0];
    bool active = false, drained = false, far = false;
    uint32_t snarl = 0, a11 = 0, a12 = 0, a21 = 0, a22 = 0;
    int rem = 0;
    double N = 0, sN = 0, D = 1, sD = 0, cur = 0, tprob = 1, cprob = 0, two = 0;
    // cur, tprob and cprob are kept as (true value) * pd0, pd0 = 2^-E.  E stays 0 unless the terms of the far side climb
    // above 2^128 (a table whose p-value is below ~1e-60): then all three are scaled down by 2^256 at the end of a block, so
    // that the climb towards the centre can go on to the reference's overflow point (cur_prob > DBL_MAX -> "0", :312-314) in
    // block form instead of overflowing the block products 250 orders early.  The result tprob / (cprob + tprob) and the
    // tests against the bias (PN_0 = cur / bias, PD_0 = pd0) do not depend on the scale.
    double pd0 = 1.0;
    // One side of the table: numerator factors n1, n2 count down, denominator factors d1, d2 count up.  The carried products
    // and their increments are scaled by k = 2^-ilogb(first denominator) — exact, and the same k on both, so every quotient
    // N_j / D_j is unchanged while the block products PD = prod D_j k stay near 1 for any M and any table size.
    auto start_side = [&](uint32_t n1, uint32_t n2, uint32_t d1, uint32_t d2) {
        const double d_first = __dmul_rn(double(d1) + 1.0, double(d2) + 1.0);
        const double k = __hiloint2double((2046 - (__double2hiint(d_first) >> 20)) << 20, 0);  // 2^-ilogb(d_first)
        N = __dmul_rn(__dmul_rn(double(n1), double(n2)), k);
        sN = __dmul_rn(double(n1) + double(n2) - 1.0, k);
        D = __dmul_rn(__dmul_rn(double(d1), double(d2)), k);
        sD = __dmul_rn(double(d1) + double(d2) + 1.0, k);
        two = __dadd_rn(k, k);
        cur = __dmul_rn(kStart, pd0);
        rem = int(n1);
    };
    auto finish = [&](double v, uint8_t code) {
        p_fisher[snarl] = v;
        fisher_code[snarl] = code;
        active = false;
    };
    for (;;) {
        // ---- refill the idle lanes from the work list ----------------------------------------------------------
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
                        fisher_big_table(t, snarl, p_fisher, fisher_code);
                    } else {
                        a11 = t.x; a12 = t.y; a21 = t.z; a22 = t.w;
                        uint32_t u;
                        if (a12 > a21) { u = a12; a12 = a21; a21 = u; }  // stats_test.cpp:280-284
                        if (a11 > a22) { u = a11; a11 = a22; a22 = u; }  // :286-290
                        if (uint64_t(a11) * a22 > uint64_t(a12) * a21) {  // :292-299
                            u = a11; a11 = a12; a12 = u;
                            u = a21; a21 = a22; a22 = u;
                        }
                        if (a12 == 0) {  // the first loop does not run (:306), cprob == 0 (:322-324)
                            p_fisher[snarl] = 1.0;
                            fisher_code[snarl] = kPOne;
                        } else {
#ifdef STOAT_FISHER_DEBUG
                            atomicAdd(&g_fisher_dbg[2], 1ull);
#endif
                            pd0 = 1.0;
                            start_side(a12, a21, a11, a22);
                            tprob = kStart; cprob = 0.0;
                            far = true;
                            active = true;
                        }
                    }
                } else {
                    drained = true;
                }
            }
            if (__any_sync(0xffffffffu, !active && !drained)) continue;  // a lane drew a table that was done at once
        }
        if (!__any_sync(0xffffffffu, active)) break;
        if (!active) {  // parked: a lane whose list ran dry iterates on harmless values
            N = 0; sN = 0; D = 1; sD = 0; two = 0; cur = 0; tprob = 1; cprob = 0; rem = 0x3FFFFFFF; far = false; pd0 = 1.0;
        }
        // ---- M steps of the recurrence, every lane in the same straight-line code ----------------------------------
        const double cur0 = cur;
        double PN = __dmul_rn(cur, kInvBias), PD = pd0, A = 0.0, T = 0.0;
#pragma unroll
        for (int j = 0; j < M; ++j) {
            D = __dadd_rn(D, sD);
            sD = __dadd_rn(sD, two);
            PN = __dmul_rn(PN, N);
            N = __dsub_rn(N, sN);
            sN = __dsub_rn(sN, two);
            PD = __dmul_rn(PD, D);
            A = fma(A, D, PN);
            // term_j < bias: the positive doubles compare like their bit patterns (a -0 past the end of the column counts as below)
            const bool below = __double_as_longlong(PN) < __double_as_longlong(PD);
            T = fma(T, D, below ? PN : 0.0);
        }
        rem -= M;
#ifdef STOAT_FISHER_DEBUG
        if (active) { atomicAdd(&g_fisher_dbg[1], 1ull); if (far) atomicAdd(&g_fisher_dbg[3], 1ull); }
#endif
        const double sc = __dmul_rn(fast_rcp(PD), __dmul_rn(kBias, pd0));
        const double t_add = __dmul_rn(T, sc), c_add = __dmul_rn(__dsub_rn(A, T), sc);
        cur = __dmul_rn(PN, sc);
        tprob = __dadd_rn(tprob, t_add);
        cprob = __dadd_rn(cprob, c_add);
        // end of the side: the last term is below the bias AND no longer counts in tprob, or the column is used up
        const double still = __dmul_rn(tprob, 1.1102230246251565e-16);
        const long long cb = __double_as_longlong(cur);
        const bool end_side = (cb < __double_as_longlong(__dmul_rn(kBias, pd0)) && cb < __double_as_longlong(still)) || rem <= 0;
        const double ovf = __dmul_rn(DBL_MAX, pd0);  // DBL_MAX on the scale of cur / tprob / cprob
        const bool nonfinite = __double2hiint(cprob) >= 0x7FF00000 || __double2hiint(tprob) >= 0x7FF00000 ||
                               __double2hiint(cur) >= 0x7FF00000;
        const bool over = c_add > ovf || cur > ovf;          // a term of this block may have passed DBL_MAX
        const bool large = __double2hiint(cur) >= 0x47F00000;  // scaled cur >= 2^128
        bool ev = (end_side || nonfinite || over || large) && active;
        if (!__any_sync(0xffffffffu, ev)) continue;
        // ---- events --------------------------------------------------------------------------------------------------
        if (ev) {
            bool passed = over && cur > ovf;  // the block's last term is beyond DBL_MAX
            if (!nonfinite && over && !passed) {
                // The block's sum passed DBL_MAX, its last term did not: whether a single term did (:312) is decided by
                // stepping through this one block again, term by term, from the state it started in (the carried products
                // are recovered exactly from the end state: they are integers times k).
                constexpr double h = double(M) * double(M - 1) / 2.0;
                double sn = __dadd_rn(sN, __dmul_rn(double(M), two)), sd = __dsub_rn(sD, __dmul_rn(double(M), two));
                double nn = __dsub_rn(__dadd_rn(N, __dmul_rn(double(M), sn)), __dmul_rn(h, two));
                double dd = __dsub_rn(__dsub_rn(D, __dmul_rn(double(M), sd)), __dmul_rn(h, two));
                double c = cur0;
                for (int j = 0; j < M && j < rem + M; ++j) {
                    dd = __dadd_rn(dd, sd);
                    sd = __dadd_rn(sd, two);
                    c = __dmul_rn(c, __ddiv_rn(nn, dd));
                    nn = __dsub_rn(nn, sn);
                    sn = __dsub_rn(sn, two);
                    if (c > ovf) {
                        passed = true;
                        break;
                    }
                }
            }
            if (nonfinite) {
                // the block products themselves left the double range (a climb of more than 2^800 within one block)
#ifdef STOAT_FISHER_DEBUG
                atomicAdd(&g_fisher_dbg[0], 1ull);
#endif
                fisher_big_table(make_uint4(a11, a12, a21, a22), snarl, p_fisher, fisher_code);
                active = false;
            } else if (passed) {
                finish(0.0, kPZero);  // :312-314: cur_prob > DBL_MAX
            } else if (!end_side) {
                // still climbing with large scaled values: scale down by 2^256 (at most three times: the bias stays a normal
                // number) and go on
                if (large && pd0 > 1e-200) {
                    const double dn = 8.636168555094445e-78;  // 2^-256
                    cur = __dmul_rn(cur, dn);
                    tprob = __dmul_rn(tprob, dn);
                    cprob = __dmul_rn(cprob, dn);
                    pd0 = __dmul_rn(pd0, dn);
                }
            } else if (far) {
                if (cprob == 0) {
                    finish(1.0, kPOne);  // :322-324: no table is more probable than the observed one
                } else if (a11 == 0) {
                    finish(__ddiv_rn(tprob, __dadd_rn(cprob, tprob)), kPValue);  // :339, :359
                } else {  // :339-344
                    start_side(a11, a22, a12, a21);
                    far = false;
                }
            } else {
                finish(__ddiv_rn(tprob, __dadd_rn(cprob, tprob)), kPValue);  // :352-354, :359
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
                                uint64_t s_count, bool beside_count_pass) {
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
    // blocks per SM: 8 when the kernel has the SMs to itself; 4 next to the persistent count pass of the following chunk
    // (416 threads x 64 registers + 4 x 128 x 64 fit one SM's register file; more would wait for the count pass to end)
    int per_sm = beside_count_pass ? k1_env("STOAT_GPU_FISHER_BLOCKS_BESIDE", 4) : k1_env("STOAT_GPU_FISHER_BLOCKS", 8);
    if (per_sm < 1) per_sm = beside_count_pass ? 4 : 8;
    const unsigned fblocks = unsigned(want < uint64_t(num_sms) * per_sm ? want : uint64_t(num_sms) * per_sm);
    // STOAT_GPU_FISHER=exact: the quotient by __ddiv_rn, bit-identical to the reference's recurrence (15 % slower)
    const char* fv = std::getenv("STOAT_GPU_FISHER");
    const bool serial_form = fv && (fv[0] == 'e' || fv[0] == 's');  // exact | serial: one step of the recurrence at a time
    if (!serial_form) {
        int per_sm_b = per_sm;
        const unsigned bb = unsigned(want < uint64_t(num_sms) * per_sm_b ? want : uint64_t(num_sms) * per_sm_b);
        // steps per block: 32 measured best on B200 (C3, stats_ms incl. K2a: 8 / 16 / 24 / 32 steps 0.59 / 0.47 / 0.41 / 0.39 ms,
        // the serial form 0.76 ms; profiles/r2g_sweep_fisher_block.json); STOAT_GPU_FISHER=8|16 select the others
        if (fv && fv[0] == '1' && fv[1] == '6')
            fisher_block_kernel<16><<<bb, 128, 0, st>>>(fisher_tab + s_begin, fisher_idx + s_begin, fisher_counters, p_fisher,
                                                        fisher_code);
        else if (fv && fv[0] == '8')
            fisher_block_kernel<8><<<bb, 128, 0, st>>>(fisher_tab + s_begin, fisher_idx + s_begin, fisher_counters, p_fisher,
                                                       fisher_code);
        else
            fisher_block_kernel<32><<<bb, 128, 0, st>>>(fisher_tab + s_begin, fisher_idx + s_begin, fisher_counters, p_fisher,
                                                        fisher_code);
#ifdef STOAT_FISHER_DEBUG
        {
            unsigned long long h[4];
            cudaStreamSynchronize(st);
            cudaMemcpyFromSymbol(h, g_fisher_dbg, sizeof(h));
            std::fprintf(stderr, "[fisher dbg] fallbacks %llu lane-blocks %llu tables %llu far-blocks %llu\n", h[0], h[1], h[2], h[3]);
            unsigned long long z[4] = {0, 0, 0, 0};
            cudaMemcpyToSymbol(g_fisher_dbg, z, sizeof(z));
        }
#endif
        return cudaGetLastError();
    }
    if (fv && fv[0] == 'e')
        fisher_exact_kernel<false><<<fblocks, 128, 0, st>>>(fisher_tab + s_begin, fisher_idx + s_begin, fisher_counters,
                                                            p_fisher, fisher_code);
    else
        fisher_exact_kernel<true><<<fblocks, 128, 0, st>>>(fisher_tab + s_begin, fisher_idx + s_begin, fisher_counters,
                                                           p_fisher, fisher_code);
    return cudaGetLastError();
}

}  // namespace stoat
