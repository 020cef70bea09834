// Micro-benchmark behind K3M's load pattern: the same bytes (64 rows per warp tile, every row streamed once) fetched as
//   A: 64-byte pieces of 8 rows per warp instruction (what K3M does: lane (g, t) reads 16 B of row g at chunk 4 * round + t)
//   B: 128-byte pieces of 4 rows per warp instruction (8 lanes per row)
//   C: 256-byte pieces of 2 rows per warp instruction (16 lanes per row)
//   D: as B with cp.async into a per-warp shared-memory ring (no registers in flight), S stages
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o k3m_pattern k3m_pattern.cu
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__device__ __forceinline__ uint4 ldg_stream(const uint4* p) {
    uint4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}
__device__ __forceinline__ uint32_t fold(uint4 v) { return v.x ^ v.y ^ v.z ^ v.w; }

constexpr int kWarps = 8;

// LANES_PER_ROW = 4 (A), 8 (B), 16 (C); a warp tile = 64 rows; a step = 64 rows x (LANES_PER_ROW * 16) bytes = 16 loads per
// lane for B (8 for A: two steps are kept in flight there to hold the same number of bytes in registers)
template <int LPR>
__global__ void __launch_bounds__(kWarps * 32, 2) pattern_ldg(const uint8_t* rows, const uint32_t* row_of, uint64_t n_tiles,
                                                               uint32_t pitch, uint32_t* out) {
    constexpr int kRowsPerInstr = 32 / LPR;
    constexpr int kLoads = 64 / kRowsPerInstr;       // loads per lane and step
    constexpr int kStepBytes = LPR * 16;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t sub = lane / LPR, c = lane % LPR;
    const uint32_t n_steps = (pitch + kStepBytes - 1) / kStepBytes;
    uint32_t acc = 0;
    for (uint64_t tile = uint64_t(blockIdx.x) * kWarps + warp; tile < n_tiles; tile += uint64_t(gridDim.x) * kWarps) {
        const uint8_t* base[kLoads];
#pragma unroll
        for (int i = 0; i < kLoads; ++i) base[i] = rows + size_t(row_of[tile * 64 + i * kRowsPerInstr + sub]) * pitch + c * 16;
        uint4 cur[kLoads], nxt[kLoads];
#pragma unroll
        for (int i = 0; i < kLoads; ++i) cur[i] = ldg_stream(reinterpret_cast<const uint4*>(base[i]));
        for (uint32_t s = 0; s < n_steps; ++s) {
            const bool more = s + 1 < n_steps && (s + 1) * kStepBytes + c * 16 < pitch;
#pragma unroll
            for (int i = 0; i < kLoads; ++i)
                nxt[i] = more ? ldg_stream(reinterpret_cast<const uint4*>(base[i] + size_t(s + 1) * kStepBytes)) : make_uint4(0, 0, 0, 0);
#pragma unroll
            for (int i = 0; i < kLoads; ++i) acc += __popc(fold(cur[i]));
#pragma unroll
            for (int i = 0; i < kLoads; ++i) cur[i] = nxt[i];
        }
    }
    if (acc == 0x12345678u) out[0] = acc;
}

// D: cp.async ring, 128 bytes per row and stage (64 rows = 8 KB per stage and warp)
template <int S>
__global__ void __launch_bounds__(kWarps * 32, 1) pattern_cpasync(const uint8_t* rows, const uint32_t* row_of, uint64_t n_tiles,
                                                                   uint32_t pitch, uint32_t* out) {
    extern __shared__ __align__(128) uint8_t smem[];
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint8_t* ring = smem + size_t(warp) * S * 8192;
    const uint32_t sub = lane >> 3, c = lane & 7;
    const uint32_t n_steps = (pitch + 127) / 128;
    uint32_t acc = 0;
    const uint32_t ring_s = uint32_t(__cvta_generic_to_shared(ring));
    for (uint64_t tile = uint64_t(blockIdx.x) * kWarps + warp; tile < n_tiles; tile += uint64_t(gridDim.x) * kWarps) {
        const uint8_t* base[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) base[i] = rows + size_t(row_of[tile * 64 + i * 4 + sub]) * pitch + c * 16;
        auto issue = [&](uint32_t s) {
            if (s < n_steps) {
                const uint32_t stage = s % S;
                const uint32_t sz = (s * 128 + c * 16 < pitch) ? 16u : 0u;
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const uint32_t r = i * 4 + sub;
                    const uint32_t dst = ring_s + stage * 8192 + r * 128 + ((c ^ (((r >> 2) & 1) << 2)) * 16);
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(base[i] + size_t(s) * 128), "r"(sz));
                }
            }
            asm volatile("cp.async.commit_group;");
        };
#pragma unroll
        for (int s = 0; s < S - 1; ++s) issue(s);
        for (uint32_t s = 0; s < n_steps; ++s) {
            issue(s + S - 1);
            asm volatile("cp.async.wait_group %0;" ::"n"(S - 1));
            __syncwarp();
            const uint8_t* st = ring + (s % S) * 8192;
            // read as K3M's lanes would: lane (g, t), rows h * 32 + g * 4 + k, chunks t and 4 + t
            const uint32_t g = lane >> 2, t = lane & 3;
#pragma unroll
            for (int half = 0; half < 2; ++half)
#pragma unroll
                for (int h = 0; h < 2; ++h)
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const uint32_t r = h * 32 + g * 4 + k;
                        const uint4 v = *reinterpret_cast<const uint4*>(st + r * 128 + (((half * 4 + t) ^ (((r >> 2) & 1) << 2)) * 16));
                        acc += __popc(fold(v));
                    }
            __syncwarp();
        }
        asm volatile("cp.async.wait_group 0;");
    }
    if (acc == 0x12345678u) out[0] = acc;
}

int main(int argc, char** argv) {
    const uint64_t n_snarls = 1000000, n_rows = n_snarls * 4;
    const uint32_t pitch = 1280;
    const int shuffle = argc > 1 ? atoi(argv[1]) : 0;
    uint8_t* rows;
    uint32_t *row_of, *out;
    CK(cudaMalloc(&rows, n_rows * pitch));
    CK(cudaMemset(rows, 0x5a, n_rows * pitch));
    CK(cudaMalloc(&out, 4));
    uint32_t* h = (uint32_t*)malloc(n_rows * 4);
    for (uint64_t i = 0; i < n_rows; ++i) h[i] = uint32_t(i);
    if (shuffle) {  // snarls keep their four rows together, the snarls themselves are permuted
        srand(1);
        for (uint64_t s = n_snarls - 1; s > 0; --s) {
            const uint64_t j = (uint64_t(rand()) * RAND_MAX + rand()) % (s + 1);
            for (int k = 0; k < 4; ++k) { uint32_t tmp = h[4 * s + k]; h[4 * s + k] = h[4 * j + k]; h[4 * j + k] = tmp; }
        }
    }
    CK(cudaMalloc(&row_of, n_rows * 4));
    CK(cudaMemcpy(row_of, h, n_rows * 4, cudaMemcpyHostToDevice));
    const uint64_t n_tiles = n_snarls / 16;
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    const double gb = double(n_rows) * pitch / 1e9;
    auto time_it = [&](const char* name, auto launch) {
        for (int i = 0; i < 3; ++i) launch();
        CK(cudaDeviceSynchronize());
        CK(cudaEventRecord(e0));
        for (int i = 0; i < 10; ++i) launch();
        CK(cudaEventRecord(e1));
        CK(cudaDeviceSynchronize());
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        printf("%-40s %.3f ms  %.0f GB/s\n", name, ms / 10, gb / (ms / 10) * 1e3);
    };
    time_it("A: 64 B x 8 rows per instr (K3M)", [&] { pattern_ldg<4><<<2 * sms, kWarps * 32>>>(rows, row_of, n_tiles, pitch, out); });
    time_it("B: 128 B x 4 rows per instr", [&] { pattern_ldg<8><<<2 * sms, kWarps * 32>>>(rows, row_of, n_tiles, pitch, out); });
    time_it("C: 256 B x 2 rows per instr", [&] { pattern_ldg<16><<<2 * sms, kWarps * 32>>>(rows, row_of, n_tiles, pitch, out); });
    CK(cudaFuncSetAttribute(pattern_cpasync<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kWarps * 2 * 8192));
    CK(cudaFuncSetAttribute(pattern_cpasync<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, kWarps * 3 * 8192));
    time_it("D2: cp.async ring, 2 stages, 8 warps/SM", [&] { pattern_cpasync<2><<<sms, kWarps * 32, kWarps * 2 * 8192>>>(rows, row_of, n_tiles, pitch, out); });
    time_it("D3: cp.async ring, 3 stages, 8 warps/SM", [&] { pattern_cpasync<3><<<sms, kWarps * 32, kWarps * 3 * 8192>>>(rows, row_of, n_tiles, pitch, out); });
    CK(cudaGetLastError());
    return 0;
}
