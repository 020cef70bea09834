// stage_pipeline.cuh — the per-warp TMA staging step shared by K1 (path_and_popc) and K3 (dosage_moments): a warp
// owns a small ring of shared-memory stages; each stage receives one snarl's path program and its distinct matrix rows
// by cp.async.bulk copies (SASS UBLKCP) that complete on the stage's mbarrier.
#pragma once
#include "ctx.hpp"
#include "ptx.cuh"

namespace stoat {

constexpr int kStageBlobBytes = kMaxBlob16 * 16;  // room for the path program at the head of a stage

// Record of a snarl as the 8 words lanes 0..7 of a warp hold: words 0..5 = rows, 6 = blob16,
// 7 = n_rows | blob_len16 << 8 | n_paths << 16.  Lane k < 6 later issues the copy of row k straight from its own
// register, lane 6 the path program, lane 7 arms the barrier: the refill costs ~10 warp instructions.
__device__ __forceinline__ uint32_t load_rec_word(const SnarlRec* rec, uint64_t s, uint32_t lane) {
    return (lane < 8) ? __ldg(reinterpret_cast<const uint32_t*>(rec + s) + lane) : 0u;
}

__device__ __forceinline__ void issue_stage(uint32_t rec_word, uint32_t lane, uint8_t* dst, uint64_t* bar,
                                            const uint8_t* rows, const uint8_t* blob, uint32_t pitch,
                                            uint64_t policy) {
    const uint32_t packed = __shfl_sync(0xffffffffu, rec_word, 7);
    const uint32_t n_rows = packed & 0xFFu, blob_bytes = ((packed >> 8) & 0xFFu) * 16u;
    // No proxy fence: the stage was only READ through the generic proxy, and those reads have completed (their
    // values were consumed, then __syncwarp) before the async proxy overwrites it — the same write-after-read
    // hand-over a consumer-release / producer-acquire mbarrier pair gives in a TMA pipeline.  (A
    // fence.proxy.async here was measured to serialise every iteration on the in-flight bulk copies.)
    if (lane == 7) mbar_arrive_expect_tx(bar, blob_bytes + n_rows * pitch);
    __syncwarp();
    // Runs of consecutive matrix rows land in consecutive row slots of the stage: one bulk copy per run (a snarl's
    // edges are usually neighbours in the matrix, so the common case is a single copy of n_rows * pitch bytes).
    const uint32_t prev = __shfl_up_sync(0xffffffffu, rec_word, 1);
    const bool head = lane < n_rows && (lane == 0 || rec_word != prev + 1u);
    const uint32_t heads = __ballot_sync(0xffffffffu, head) | (1u << n_rows);
    if (head) {
        const uint32_t run = uint32_t(__ffs(int(heads >> (lane + 1))));  // distance to the next run (or the end)
        tma_bulk_g2s_hint(dst + kStageBlobBytes + lane * pitch, rows + size_t(rec_word) * pitch, run * pitch, bar, policy);
    } else if (lane == 6) {
        tma_bulk_g2s(dst, blob + size_t(rec_word) * 16, blob_bytes, bar);
    }
}

// Stage refill for a slice of the haplotype range: like issue_stage (stage_pipeline.cuh), but every row contributes only
// the bytes [slice_off, slice_off + slice_bytes) and lands in a slot of slot_bytes.  Whole rows (slice_bytes == pitch):
// runs of consecutive matrix rows still go out as one bulk copy.
__device__ __forceinline__ void issue_stage_slice(uint32_t rec_word, uint32_t lane, uint8_t* dst, uint64_t* bar,
                                                  const uint8_t* rows, const uint8_t* blob, uint32_t pitch,
                                                  uint32_t slice_off, uint32_t slice_bytes, uint32_t slot_bytes,
                                                  uint64_t policy) {
    const uint32_t packed = __shfl_sync(0xffffffffu, rec_word, 7);
    const uint32_t n_rows = packed & 0xFFu, blob_bytes = ((packed >> 8) & 0xFFu) * 16u;
    if (lane == 7) mbar_arrive_expect_tx(bar, blob_bytes + n_rows * slice_bytes);
    __syncwarp();
    const bool whole = slice_bytes == pitch && slot_bytes == pitch;
    const uint32_t prev = __shfl_up_sync(0xffffffffu, rec_word, 1);
    const bool head = lane < n_rows && (!whole || lane == 0 || rec_word != prev + 1u);
    const uint32_t heads = __ballot_sync(0xffffffffu, head) | (1u << n_rows);
    if (head) {
        const uint32_t run = whole ? uint32_t(__ffs(int(heads >> (lane + 1)))) : 1u;
        tma_bulk_g2s_hint(dst + kStageBlobBytes + lane * slot_bytes, rows + size_t(rec_word) * pitch + slice_off,
                          run * slice_bytes, bar, policy);
    } else if (lane == 6) {
        tma_bulk_g2s(dst, blob + size_t(rec_word) * 16, blob_bytes, bar);
    }
}

}  // namespace stoat
