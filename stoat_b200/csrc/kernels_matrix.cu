// kernels_matrix.cu — K0 repack_rows and the phenotype bit mask.
//
// The reference stores the presence matrix as one unaligned bit stream: bit (row * n_cols + col),
// LSB-first (src/matrix.cpp:67-84).  With n_cols = 10,000 a row is 1,250 bytes and starts at an
// arbitrary bit, so neither 128-bit loads nor TMA bulk copies can address it.  K0 rewrites the
// stream once into rows of `pitch` = ceil(n_cols/128)*16 bytes, 16-byte aligned, padding bits zero.
// HBM-bound: reads n_rows*n_cols/8 bytes, writes n_rows*pitch bytes.
#include "ctx.hpp"
#include "ptx.cuh"

namespace stoat {

__device__ __forceinline__ uint32_t load_word_guarded(const uint8_t* raw, uint64_t raw_bytes, uint64_t word_idx) {
    const uint64_t byte0 = word_idx * 4;
    if (byte0 + 4 <= raw_bytes) return __ldg(reinterpret_cast<const uint32_t*>(raw) + word_idx);
    uint32_t v = 0;
    for (int k = 0; k < 4; ++k)
        if (byte0 + k < raw_bytes) v |= uint32_t(__ldg(raw + byte0 + k)) << (8 * k);
    return v;
}

// one warp per row, lanes stride over the row's 16-byte chunks
__global__ void __launch_bounds__(256) repack_rows_kernel(const uint8_t* __restrict__ raw, uint64_t raw_bytes,
                                                          uint64_t row_begin, uint64_t row_end, uint64_t n_cols,
                                                          uint32_t chunks_per_row, uint8_t* __restrict__ out) {
    const uint32_t lane = threadIdx.x & 31;
    const uint64_t warp = (uint64_t(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const uint64_t n_warps = (uint64_t(gridDim.x) * blockDim.x) >> 5;
    for (uint64_t row = row_begin + warp; row < row_end; row += n_warps) {
        const uint64_t bit_base = row * n_cols;
        uint4* dst = reinterpret_cast<uint4*>(out + row * uint64_t(chunks_per_row) * 16);
        for (uint32_t c = lane; c < chunks_per_row; c += 32) {
            const uint64_t b0 = bit_base + uint64_t(c) * 128;
            const uint64_t w0 = b0 >> 5;
            const uint32_t sh = uint32_t(b0 & 31);
            uint32_t w[5];
#pragma unroll
            for (int k = 0; k < 5; ++k) w[k] = load_word_guarded(raw, raw_bytes, w0 + k);
            uint32_t o[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                o[k] = __funnelshift_r(w[k], w[k + 1], sh);
                // zero the padding: keep only columns < n_cols
                const uint64_t col0 = uint64_t(c) * 128 + 32 * k;
                if (col0 >= n_cols)
                    o[k] = 0;
                else if (col0 + 32 > n_cols)
                    o[k] &= (1u << uint32_t(n_cols - col0)) - 1u;
            }
            stg_stream(dst + c, make_uint4(o[0], o[1], o[2], o[3]));
        }
    }
}

cudaError_t launch_repack_rows(const uint8_t* raw, uint64_t raw_bytes, uint64_t row_begin, uint64_t row_end,
                               uint64_t n_cols, uint64_t pitch, uint8_t* out, cudaStream_t st) {
    if (row_end <= row_begin) return cudaSuccess;
    const uint64_t n = row_end - row_begin;
    const uint64_t warps_per_block = 8;
    uint64_t blocks = (n + warps_per_block - 1) / warps_per_block;
    const uint64_t max_blocks = 148ull * 8 * 16;  // grid-stride beyond this
    if (blocks > max_blocks) blocks = max_blocks;
    repack_rows_kernel<<<unsigned(blocks), 256, 0, st>>>(raw, raw_bytes, row_begin, row_end, n_cols,
                                                         uint32_t(pitch / 16), out);
    return cudaGetLastError();
}

// bit h of the mask = is_case[h / 2] for h < 2 * n_samples; rest of the pitch zero
__global__ void build_case_mask_kernel(const uint8_t* __restrict__ is_case, uint64_t n_samples, uint32_t n_words,
                                       uint32_t* __restrict__ mask) {
    const uint32_t w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= n_words) return;
    uint32_t v = 0;
    const uint64_t s0 = uint64_t(w) * 16;  // 16 samples = 32 haplotypes per word
    for (int k = 0; k < 16; ++k) {
        const uint64_t s = s0 + k;
        if (s < n_samples && is_case[s]) v |= 3u << (2 * k);
    }
    mask[w] = v;
}

cudaError_t launch_build_case_mask(const uint8_t* is_case, uint64_t n_samples, uint64_t pitch, uint8_t* mask,
                                   cudaStream_t st) {
    const uint32_t n_words = uint32_t(pitch / 4);
    if (n_words == 0) return cudaSuccess;
    build_case_mask_kernel<<<(n_words + 255) / 256, 256, 0, st>>>(is_case, n_samples, n_words,
                                                                  reinterpret_cast<uint32_t*>(mask));
    return cudaGetLastError();
}

// =================================================================================================
// Matrix from its set bits — the device half of make_edge_matrix (src/snarl_analyzer.cpp:190-260): while it parses a VCF
// record the reference sets bit (row of the allele edge, 2 * sample + haplotype) one at a time (matrix.cpp:76-84); here
// the host hands over those (row, column) pairs and the bits are set straight into the pitched layout (no K0, and the
// upload is 8 bytes per set bit instead of the dense matrix: smaller below 1.5 % density).  `out` is zeroed by the caller.
// =================================================================================================
__global__ void __launch_bounds__(256) scatter_bits_kernel(const uint32_t* __restrict__ row, const uint32_t* __restrict__ col,
                                                           uint64_t n_pairs, uint64_t n_rows, uint64_t n_cols, uint64_t pitch,
                                                           uint32_t* __restrict__ out, uint32_t* __restrict__ bad) {
    const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
    for (uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n_pairs; i += stride) {
        const uint32_t r = row[i], c = col[i];
        if (uint64_t(r) >= n_rows || uint64_t(c) >= n_cols) {
            atomicExch(bad, 1u);
            continue;
        }
        atomicOr(out + uint64_t(r) * (pitch / 4) + (c >> 5), 1u << (c & 31u));
    }
}

cudaError_t launch_scatter_bits(const uint32_t* row, const uint32_t* col, uint64_t n_pairs, uint64_t n_rows, uint64_t n_cols,
                                uint64_t pitch, uint8_t* out, uint32_t* bad, int num_sms, cudaStream_t st) {
    if (n_pairs == 0) return cudaSuccess;
    const uint64_t want = (n_pairs + 255) / 256;
    const unsigned blocks = unsigned(want < uint64_t(num_sms) * 16 ? want : uint64_t(num_sms) * 16);
    scatter_bits_kernel<<<blocks, 256, 0, st>>>(row, col, n_pairs, n_rows, n_cols, pitch, reinterpret_cast<uint32_t*>(out),
                                                bad);
    return cudaGetLastError();
}


// rows[i] of the pitched matrix -> out[i] (same pitch): the row slices a sample-sharded run exchanges for the snarls its
// summed two-path partials cannot finish (stoat_b200/shard.py).  One block per row, 16-byte copies.
__global__ void __launch_bounds__(128) gather_rows_kernel(const uint8_t* __restrict__ rows, uint64_t pitch,
                                                          const uint32_t* __restrict__ idx, uint8_t* __restrict__ out) {
    const uint4* src = reinterpret_cast<const uint4*>(rows + uint64_t(idx[blockIdx.x]) * pitch);
    uint4* dst = reinterpret_cast<uint4*>(out + uint64_t(blockIdx.x) * pitch);
    for (uint32_t c = threadIdx.x; c < pitch / 16; c += blockDim.x) dst[c] = src[c];
}

cudaError_t launch_gather_rows(const uint8_t* rows, uint64_t pitch, const uint32_t* idx, uint64_t n, uint8_t* out,
                               cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    gather_rows_kernel<<<unsigned(n), 128, 0, st>>>(rows, pitch, idx, out);
    return cudaGetLastError();
}

}  // namespace stoat
