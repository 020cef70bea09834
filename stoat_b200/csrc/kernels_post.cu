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
