// kernels_post.cu — Benjamini-Hochberg adjustment of the p-value column on the device.
//
//   adjust_pvalues_with_BH (src/post_processing.cpp:7-38): sort by raw p; adjusted[i] = p * n / (i + 1); running minimum
//   from the largest p down; clamp to 1; back to the original order.
// Here: radix sort of (p, index) (non-negative doubles order like their bit patterns), one elementwise kernel with the
// reference's two roundings (`__dmul_rn`, `__ddiv_rn`), a reverse inclusive min-scan (exact and associative, so the scan
// order does not matter) and a scatter.  Tied p-values get the same adjusted value whatever order the sort leaves them in.
// The sort and the scan are CUB device primitives (library code): this is post-processing, not the hot path.
#include <cub/cub.cuh>

#include <mutex>

#include "ctx.hpp"
#include "device_math.cuh"

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

// ---- Benjamini-Hochberg without any allocation (stoat_gpu_bh_adjust_static) ---------------------------------------
// The same kernels as launch_bh_adjust with their scratch (sort keys, index pairs, the two scan arrays, CUB's temporary
// storage, one flag word) in a file-scope device array instead of a context-owned cudaMalloc block, for callers that
// cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  Every GPU that loads the
// library holds kBhStaticBytes (64 MiB) until the process exits.  The array is shared by every call and every context of
// the process on that GPU and keeps the contents of the last call: calls that overlap on the GPU race on it.
constexpr uint64_t kBhStaticMaxN = uint64_t(STOAT_GPU_BH_STATIC_MAX_N);
constexpr size_t kBhStaticFixed = size_t(kBhStaticMaxN) * (5 * 8 + 2 * 4);  // p | out | keys | a | b | idx_in | idx_sorted
constexpr size_t kBhStaticTemp = size_t(16) << 20;                           // CUB temporary storage (checked at set-up)
constexpr size_t kBhStaticBytes = kBhStaticFixed + kBhStaticTemp;
static __device__ __align__(256) uint8_t g_bh_static_scratch[kBhStaticBytes];

// 1 in *flag when some p[i] is negative, NaN or infinite (the host check of stoat_gpu_bh_adjust, done where the data is)
__global__ void __launch_bounds__(256) bh_check_kernel(const double* __restrict__ p, uint64_t n, uint32_t* flag) {
    const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < n && (!(p[i] >= 0.0) || isinf(p[i]))) atomicOr(flag, 1u);
}

// Device address of the array on `device`, resolved once per device with that device current.  In host code the array's
// NAME is only a host-side stand-in for the symbol: a kernel must be given the address cudaGetSymbolAddress returns.
static cudaError_t bh_static_base(int device, uint8_t** base) {
    constexpr int kMaxDev = 64;
    static std::mutex mu;
    static uint8_t* addr[kMaxDev] = {};
    if (device < 0 || device >= kMaxDev) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(mu);
    if (!addr[device]) {
        int cur = -1;
        cudaError_t e = cudaGetDevice(&cur);
        if (e != cudaSuccess) return e;
        if (cur != device) return cudaErrorInvalidDevice;  // the caller makes the device current first
        void* ptr = nullptr;
        e = cudaGetSymbolAddress(&ptr, g_bh_static_scratch);
        if (e != cudaSuccess) return e;
        if (bh_temp_bytes(kBhStaticMaxN) > kBhStaticTemp) return cudaErrorMemoryAllocation;  // CUB wants more than reserved
        addr[device] = static_cast<uint8_t*>(ptr);
    }
    *base = addr[device];
    return cudaSuccess;
}

uint64_t bh_static_max_n() { return kBhStaticMaxN; }

// p / out: host or device pointers of the caller (unified addressing tells which), n <= kBhStaticMaxN.  *bad = 1 when the
// input held a value the reference's string_to_pvalue would not have produced (nothing is written to out then).
cudaError_t launch_bh_adjust_static(int device, const double* p, uint64_t n, double* out, uint32_t* bad, cudaStream_t st) {
    *bad = 0;
    if (n == 0) return cudaSuccess;
    if (n > kBhStaticMaxN) return cudaErrorInvalidValue;
    uint8_t* base = nullptr;
    cudaError_t e = bh_static_base(device, &base);
    if (e != cudaSuccess) return e;
    double* d_p = reinterpret_cast<double*>(base);
    double* d_out = d_p + kBhStaticMaxN;
    uint64_t* d_keys = reinterpret_cast<uint64_t*>(d_out + kBhStaticMaxN);
    double* d_a = reinterpret_cast<double*>(d_keys + kBhStaticMaxN);
    double* d_b = d_a + kBhStaticMaxN;
    uint32_t* d_i0 = reinterpret_cast<uint32_t*>(d_b + kBhStaticMaxN);
    uint32_t* d_i1 = d_i0 + kBhStaticMaxN;
    uint8_t* d_temp = base + kBhStaticFixed;
    // the flag word sits at the end of the temporary storage (bh_temp_bytes leaves 256 spare bytes) and must start at 0:
    // the array keeps whatever the last call left in it
    uint32_t* d_flag = reinterpret_cast<uint32_t*>(d_temp + kBhStaticTemp - 256);
    if ((e = cudaMemsetAsync(d_flag, 0, 4, st)) != cudaSuccess) return e;
    if ((e = cudaMemcpyAsync(d_p, p, n * 8, cudaMemcpyDefault, st)) != cudaSuccess) return e;
    bh_check_kernel<<<unsigned((n + 255) / 256), 256, 0, st>>>(d_p, n, d_flag);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    if ((e = cudaMemcpyAsync(bad, d_flag, 4, cudaMemcpyDeviceToHost, st)) != cudaSuccess) return e;
    if ((e = cudaStreamSynchronize(st)) != cudaSuccess) return e;
    if (*bad) return cudaSuccess;
    e = launch_bh_adjust(d_p, n, d_out, d_keys, d_i0, d_i1, d_a, d_b, d_temp, kBhStaticTemp - 256, st);
    if (e != cudaSuccess) return e;
    if ((e = cudaMemcpyAsync(out, d_out, n * 8, cudaMemcpyDefault, st)) != cudaSuccess) return e;
    return cudaStreamSynchronize(st);
}


// =================================================================================================
// Output formatting on the device: the value columns of the TSV rows as text.
//
//   stoat::set_precision (src/utils.cpp:5-15): |v| < 0.1 and v != 0 -> iostream scientific, precision 4 ("%.4e"), else
//   defaultfloat, precision 4 ("%.4g").  The reference builds an ostringstream per value (and flushes the file per row,
//   src/writer.cpp:34,49,85); here one thread turns one value into a 16-byte slot [text, NUL padded] + its length.
// The decimal digits come from v * 10^k rounded to an integer in FP64, which is the correctly rounded decimal unless
// v * 10^k lies within 1e-6 of a rounding boundary; those values (about one in 10^5) are flagged and the caller formats
// them with the C library.  So every string is the reference's string, and the host only assembles rows.
// =================================================================================================
__constant__ double kPow10Small[23] = {1e0,  1e1,  1e2,  1e3,  1e4,  1e5,  1e6,  1e7,  1e8,  1e9,  1e10, 1e11,
                                       1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};
__constant__ double kPow10Big[16] = {1e0,   1e22,  1e44,  1e66,  1e88,  1e110, 1e132, 1e154,
                                     1e176, 1e198, 1e220, 1e242, 1e264, 1e286, 1e308, 1e308};

// a * 10^k for |k| <= 345 without leaving the double range on the way (a is finite and positive)
__device__ __forceinline__ double scale_pow10(double a, int k) {
    const int kk = k < 0 ? -k : k;
    int big = kk / 22, small = kk - 22 * big;
    double r = a;
    if (k >= 0) {
        r *= kPow10Small[small];
        if (big > 14) { r *= 1e308; big -= 14; }
        r *= kPow10Big[big];
    } else {
        r /= kPow10Small[small];
        if (big > 14) { r /= 1e308; big -= 14; }
        r /= kPow10Big[big];
    }
    return r;
}

__device__ __forceinline__ int put_uint(char* o, uint32_t v, int width) {  // `width` digits, zero padded
    for (int i = width - 1; i >= 0; --i) {
        o[i] = char('0' + v % 10u);
        v /= 10u;
    }
    return width;
}

// returns the length; `unsure` = the decimal rounding was too close to call in FP64
__device__ int format_like_set_precision(double v, char* o, bool& unsure) {
    unsure = false;
    int n = 0;
    if (isnan(v)) {
        if (signbit(v)) o[n++] = '-';
        o[n++] = 'n'; o[n++] = 'a'; o[n++] = 'n';
        return n;
    }
    if (signbit(v)) o[n++] = '-';
    const double a = fabs(v);
    if (isinf(a)) {
        o[n++] = 'i'; o[n++] = 'n'; o[n++] = 'f';
        return n;
    }
    if (a == 0.0) {
        o[n++] = '0';
        return n;
    }
    const bool sci5 = a < 1e-1;             // "%.4e": 5 significant digits
    const int nd = sci5 ? 5 : 4;            // "%.4g": 4 significant digits
    int e10 = int(floor(log10(a)));
    const double lo = sci5 ? 1e4 : 1e3, hi = sci5 ? 1e5 : 1e4;
    double x = scale_pow10(a, nd - 1 - e10);
    if (x < lo) { --e10; x = scale_pow10(a, nd - 1 - e10); }
    if (x >= hi) { ++e10; x = scale_pow10(a, nd - 1 - e10); }
    const double fl = floor(x);
    unsure = fabs((x - fl) - 0.5) < 1e-6 || x - lo < 1e-6 || hi - x < 1e-6;
    uint32_t m = uint32_t(rint(x));
    if (m >= uint32_t(hi)) {  // 9.99996 -> 10.000
        m = uint32_t(lo);
        ++e10;
    }
    if (sci5 || e10 < -4 || e10 >= 4) {
        // d.dddd e+-XX ; "%g" drops trailing zeros (and a bare point)
        char dg[5];
        put_uint(dg, m, nd);
        int last = nd - 1;
        if (!sci5)
            while (last > 0 && dg[last] == '0') --last;
        o[n++] = dg[0];
        if (last > 0) {
            o[n++] = '.';
            for (int i = 1; i <= last; ++i) o[n++] = dg[i];
        }
        o[n++] = 'e';
        o[n++] = e10 < 0 ? '-' : '+';
        const uint32_t ea = uint32_t(e10 < 0 ? -e10 : e10);
        n += put_uint(o + n, ea, ea >= 100 ? 3 : 2);
        return n;
    }
    // "%g", fixed notation: 4 significant digits, trailing zeros dropped.  e10 in [-4, 3]
    char dg[4];
    put_uint(dg, m, 4);
    int last = 3;
    while (last > 0 && dg[last] == '0') --last;
    if (e10 >= 0) {
        for (int i = 0; i <= e10; ++i) o[n++] = dg[i];  // integer part (digits beyond `last` are zeros that stay)
        if (last > e10) {
            o[n++] = '.';
            for (int i = e10 + 1; i <= last; ++i) o[n++] = dg[i];
        }
    } else {
        o[n++] = '0';
        o[n++] = '.';
        for (int i = 0; i < -e10 - 1; ++i) o[n++] = '0';
        for (int i = 0; i <= last; ++i) o[n++] = dg[i];
    }
    return n;
}

__global__ void __launch_bounds__(256) format_values_kernel(const double* __restrict__ v, const uint8_t* __restrict__ code,
                                                            uint64_t n, char* __restrict__ slots, uint8_t* __restrict__ len,
                                                            uint8_t* __restrict__ unsure_out) {
    const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    char o[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) o[k] = 0;
    int l = 0;
    bool unsure = false;
    const uint8_t c = code ? code[i] : uint8_t(kPValue);
    if (c == kPNa) {
        o[0] = 'N'; o[1] = 'A'; l = 2;
    } else if (c == kPZero) {
        o[0] = '0'; l = 1;
    } else if (c == kPOne) {
        o[0] = '1'; l = 1;
    } else {
        l = format_like_set_precision(c == kPDblMax ? DBL_MAX : v[i], o, unsure);
    }
    uint4 w;
    memcpy(&w, o, 16);
    reinterpret_cast<uint4*>(slots)[i] = w;
    len[i] = uint8_t(l);
    unsure_out[i] = unsure ? 1 : 0;
}

cudaError_t launch_format_values(const double* v, const uint8_t* code, uint64_t n, char* slots, uint8_t* len,
                                 uint8_t* unsure, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    format_values_kernel<<<unsigned((n + 255) / 256), 256, 0, st>>>(v, code, n, slots, len, unsure);
    return cudaGetLastError();
}

}  // namespace stoat
