// kernels_post.cu — Benjamini-Hochberg adjustment of the p-value column on the device.
//
//   adjust_pvalues_with_BH (src/post_processing.cpp:7-38): sort by raw p; adjusted[i] = p * n / (i + 1); running minimum
//   from the largest p down; clamp to 1; back to the original order.
// Here: radix sort of (p, index) (non-negative doubles order like their bit patterns), one elementwise kernel with the
// reference's two roundings (`__dmul_rn`, `__ddiv_rn`), a reverse inclusive min-scan (exact and associative, so the scan
// order does not matter) and a scatter.  Tied p-values get the same adjusted value whatever order the sort leaves them in.
// The sort and the scan are CUB device primitives (library code): this is post-processing, not the hot path.
#include <cub/cub.cuh>

#include "ctx.hpp"

namespace stoat {

__global__ void __launch_bounds__(256) bh_iota_kernel(uint32_t* idx, uint64_t n) {
    const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < n) idx[i] = uint32_t(i);
}

// adjusted[i] = p_sorted[i] * n / (i + 1), stored REVERSED (largest p first) so that a forward min-scan is the reference's
// backward pass (:26-28)
__global__ void __launch_bounds__(256) bh_scale_kernel(const double* __restrict__ p_sorted, uint64_t n,
                                                       double* __restrict__ adj_rev) {
    const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    adj_rev[n - 1 - i] = __ddiv_rn(__dmul_rn(p_sorted[i], double(n)), double(i + 1));  // :21-22
}

__global__ void __launch_bounds__(256) bh_scatter_kernel(const double* __restrict__ cummin_rev,
                                                         const uint32_t* __restrict__ idx_sorted, uint64_t n,
                                                         double* __restrict__ out) {
    const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[idx_sorted[i]] = fmin(1.0, cummin_rev[n - 1 - i]);  // :31-33, original order (:36-38)
}

struct MinOp {
    __device__ __forceinline__ double operator()(double a, double b) const { return b < a ? b : a; }
};

size_t bh_temp_bytes(uint64_t n) {
    size_t a = 0, b = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, a, static_cast<const uint64_t*>(nullptr), static_cast<uint64_t*>(nullptr),
                                    static_cast<const uint32_t*>(nullptr), static_cast<uint32_t*>(nullptr), int(n));
    cub::DeviceScan::InclusiveScan(nullptr, b, static_cast<const double*>(nullptr), static_cast<double*>(nullptr), MinOp(),
                                   int(n));
    return (a > b ? a : b) + 256;
}

// p (device, [n], finite and >= 0), out (device, [n]); scratch: keys [n] u64, idx_in / idx_out [n] u32, a / b [n] f64, temp
cudaError_t launch_bh_adjust(const double* p, uint64_t n, double* out, uint64_t* keys_sorted, uint32_t* idx_in,
                             uint32_t* idx_sorted, double* a, double* b, void* temp, size_t temp_bytes, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    const unsigned blocks = unsigned((n + 255) / 256);
    bh_iota_kernel<<<blocks, 256, 0, st>>>(idx_in, n);
    cudaError_t e = cub::DeviceRadixSort::SortPairs(temp, temp_bytes, reinterpret_cast<const uint64_t*>(p), keys_sorted,
                                                    idx_in, idx_sorted, int(n), 0, 64, st);
    if (e != cudaSuccess) return e;
    bh_scale_kernel<<<blocks, 256, 0, st>>>(reinterpret_cast<const double*>(keys_sorted), n, a);
    e = cub::DeviceScan::InclusiveScan(temp, temp_bytes, a, b, MinOp(), int(n), st);
    if (e != cudaSuccess) return e;
    bh_scatter_kernel<<<blocks, 256, 0, st>>>(b, idx_sorted, n, out);
    return cudaGetLastError();
}

}  // namespace stoat
