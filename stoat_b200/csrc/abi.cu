// abi.cu — implementation of the C ABI declared in include/stoat_gpu.h.
//
// Host side only: argument checks, device buffers, copies, kernel launches, timing with CUDA events
// on the context's stream.  No statistics are computed on the host: if the device is not a usable
// sm_100 GPU every entry point fails (there is no CPU fallback).
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <thread>
#include <vector>

#include "ctx.hpp"
#include "eqtl_dense.hpp"

using namespace stoat;

namespace {

thread_local std::string g_create_error;

int fail(stoat_gpu_ctx* ctx, int code, const std::string& msg) {
    if (ctx)
        ctx->err = msg;
    else
        g_create_error = msg;
    return code;
}
int fail_cuda(stoat_gpu_ctx* ctx, cudaError_t e, const char* where) {
    const int code = (e == cudaErrorMemoryAllocation) ? STOAT_GPU_ERR_OOM : STOAT_GPU_ERR_CUDA;
    return fail(ctx, code, std::string(where) + ": " + cudaGetErrorString(e));
}
#define CK(expr)                                                  \
    do {                                                          \
        cudaError_t e__ = (expr);                                 \
        if (e__ != cudaSuccess) return fail_cuda(ctx, e__, #expr); \
    } while (0)

// Row pitch: whole 32-byte sectors.  K3M's lane groups read 64 bytes of a row at a time; with 16-byte aligned rows every
// second row made those reads straddle three sectors (ncu: 26 % more DRAM traffic than the algorithmic bytes).
uint64_t pitch_for(uint64_t n_cols) { return ((n_cols + 255) / 256) * 32; }

MatrixDev matrix_view(const stoat_gpu_ctx* ctx) {
    MatrixDev m;
    m.rows = ctx->rows.as<uint8_t>();
    m.n_rows = ctx->n_rows;
    m.n_cols = ctx->n_cols;
    m.pitch = ctx->pitch;
    return m;
}

struct QuantState {
    DevBuf yc, ys, mom, part_i, part_d, slow, yslices, lean_rec, tail_t, tail_x, tail_df;
    DevBuf cov_flag;             // eQTL with covariates: a byte per snarl of a launch_eqtl_general call
    DevBuf cov_mom, cov_prefix;  // covariate runs (quant_covar.cuh): per-snarl moments, prefix sums of the covariate rows
    double sy_all = 0.0, syy_all = 0.0, y_scale = 1.0;
    bool have_partials = false;
    uint64_t lean_key = 0;  // what lean_rec holds: staging generation + participating kernels (launch_quantitative_impl)
};

struct LogitState {  // logistic regression (binary trait with covariates): its own result buffers
    DevBuf counts, list, n_list, allele, filtered, n_used, p, beta, se, code;
    bool have_result = false;
    uint32_t n_general = 0;
};

}  // namespace

constexpr int kHostRing = 3;
constexpr uint64_t kHostRingBytes = 64ull << 20;
// memcpy by a few threads: one core moves ~10 GB/s, the PCIe link 55
static void parallel_memcpy(uint8_t* dst, const uint8_t* src, uint64_t n) {
    static const unsigned n_threads = [] {
        unsigned hw = std::thread::hardware_concurrency();
        if (const char* e = std::getenv("STOAT_GPU_UPLOAD_THREADS")) hw = unsigned(std::max(1, std::atoi(e))) * 2;
        return std::max(1u, std::min(6u, hw / 2));
    }();
    if (n_threads == 1 || n < (4ull << 20)) {
        std::memcpy(dst, src, n);
        return;
    }
    const uint64_t per = ((n + n_threads - 1) / n_threads + 4095) & ~uint64_t(4095);
    std::vector<std::thread> th;
    for (unsigned t = 1; t < n_threads; ++t) {
        const uint64_t a = std::min(n, per * t), b = std::min(n, per * (t + 1));
        if (b > a) th.emplace_back([=] { std::memcpy(dst + a, src + a, b - a); });
    }
    std::memcpy(dst, src, std::min(n, per));
    for (std::thread& t : th) t.join();
}
constexpr int kMaxRunChunks = 8;  // chunks of an overlapped binary run (stoat_gpu_binary_run_staged)
// Chunks of the binary run by default.  Measured on B200 at C3 with the 64-register count pass and the branch-free Fisher loop
// (profiles/r2a_sweep_c3.json): 1 chunk 1.57 ms, 2 / 4 / 8 chunks 1.80 / 1.90 / 2.10 ms (4 chunks with 8 Fisher blocks per
// SM: 1.64 ms).  Both kernels fit one SM now, but both are bound by the SM's issue slots (K1 by its shared-memory loads and
// popcounts, the Fisher list by the FP64 pipe), so running them side by side only stretches both: back to back stays.
constexpr int kDefaultRunChunks = 1;
constexpr int kRunRing = 64;      // staged runs whose events are kept for stoat_gpu_mean_timing

// extra per-context state that does not need to be visible to the kernel translation units
struct CtxExtra {
    QuantState q;
    LogitState lg;
    uint32_t max_stage_rows = 0;
    uint64_t stage_gen = 0;  // bumped by every stage_paths
    uint32_t counters_host[4] = {0, 0, 0, 0};
    uint32_t readback[2] = {0, 0};
    uint32_t fisher_readback[2 * kMaxRunChunks] = {};
    DevBuf bh;  // scratch of stoat_gpu_bh_adjust
    void* host_ring[kHostRing] = {};  // pinned staging of a pageable matrix upload (set_matrix_common)
    cudaEvent_t host_ring_free[kHostRing] = {};
    DevBuf um_scale, um_b, um_a, um_probe;  // tcgen05 eQTL: gene scales, B / A operand images, probe scratch
    DevBuf fmt;  // scratch of stoat_gpu_format_values
    DevBuf gather_idx, gather_out;  // stoat_gpu_gather_rows_dev
    std::vector<uint8_t> fmt_unsure;
    cudaEvent_t chunk_done[kMaxRunChunks] = {};
    // Events of the last kRunRing staged runs: [0] start, [1] end of the count / moments pass, [2] start and [3] end of the
    // statistics.  A run only enqueues work; its timing (and the counters that come back with it) is read when somebody asks
    // (finalize_timing), so that back-to-back runs keep the GPU busy without a host round trip in between.
    cudaEvent_t ring[kRunRing][4] = {};
    uint8_t ring_kind[kRunRing] = {};  // 1 binary, 2 quantitative
    uint64_t run_seq = 0;
    int pending = 0;         // kind of the run whose timing has not been read yet (0: none)
    int pending_chunks = 1;  // chunks of that (binary) run
    uint64_t expr_cols = 0;  // n_cols the expression matrix was centred and padded for
};
static CtxExtra* extra_of(stoat_gpu_ctx* ctx);

struct stoat_gpu_ctx_full : stoat_gpu_ctx {
    CtxExtra extra;
};
static CtxExtra* extra_of(stoat_gpu_ctx* ctx) { return &static_cast<stoat_gpu_ctx_full*>(ctx)->extra; }

// Read the timing of the last staged run (and the counters that belong to it); the run itself only enqueued work.
static int finalize_timing(stoat_gpu_ctx* ctx) {
    CtxExtra* x = extra_of(ctx);
    if (!x->pending) return STOAT_GPU_OK;
    const int kind = x->pending;
    x->pending = 0;
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    cudaEvent_t* ev = x->ring[(x->run_seq - 1) % kRunRing];
    const uint64_t S = ctx->batch.n_snarls;
    if (kind == 1) {
        CK(cudaMemcpy(x->fisher_readback, ctx->fisher_cnt.p, 8 * kMaxRunChunks, cudaMemcpyDeviceToHost));
        uint32_t n_fisher = 0;
        for (int c = 0; c < x->pending_chunks; ++c) n_fisher += x->fisher_readback[2 * c];
        ctx->timing.n_general = ctx->n_big;
        ctx->timing.n_fisher = S ? n_fisher : 0;
        CK(cudaEventElapsedTime(&ctx->timing.table_ms, ev[0], ev[1]));
        CK(cudaEventElapsedTime(&ctx->timing.stats_ms, ev[2], ev[3]));
    } else {
        CK(cudaMemcpy(x->readback, ctx->d_counters.p, 4, cudaMemcpyDeviceToHost));
        ctx->timing.n_general = S ? x->readback[0] : 0;
        ctx->timing.n_fisher = 0;
        if (S && x->readback[0] && std::getenv("STOAT_GPU_DEBUG_GENERAL")) {  // which snarls took the general path
            std::vector<uint32_t> ids(std::min<uint32_t>(x->readback[0], 64));
            CK(cudaMemcpy(ids.data(), ctx->batch.big_list, ids.size() * 4, cudaMemcpyDeviceToHost));
            std::fprintf(stderr, "[stoat_gpu] general path: %u snarls:", x->readback[0]);
            for (uint32_t id : ids) std::fprintf(stderr, " %u", id);
            std::fprintf(stderr, "\n");
        }
        if (S) {
            CK(cudaEventElapsedTime(&ctx->timing.table_ms, ev[0], ev[1]));  // the moments pass (K3M / K3R / K3)
            CK(cudaEventElapsedTime(&ctx->timing.stats_ms, ev[1], ev[3]));  // left-over list, K4, general path
        } else {
            ctx->timing.table_ms = ctx->timing.stats_ms = 0.f;
        }
    }
    return STOAT_GPU_OK;
}

// a new matrix invalidates everything that was derived from the old one
static void invalidate_derived(stoat_gpu_ctx* ctx) {
    ctx->have_batch = false;
    ctx->have_expression = ctx->have_eqtl_summary = false;
    ctx->have_binary_counts = ctx->have_binary_result = ctx->have_quant_result = false;
    extra_of(ctx)->q.have_partials = false;
    extra_of(ctx)->q.lean_key = 0;
    extra_of(ctx)->lg.have_result = false;
}

extern "C" {

int stoat_gpu_abi_version(void) { return STOAT_GPU_ABI_VERSION; }

const char* stoat_gpu_last_error(const stoat_gpu_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

int stoat_gpu_create(stoat_gpu_ctx** out, int device) {
    stoat_gpu_ctx* ctx = nullptr;
    if (!out) return fail(nullptr, STOAT_GPU_ERR_ARG, "stoat_gpu_create: out is NULL");
    *out = nullptr;
    int n_dev = 0;
    cudaError_t e = cudaGetDeviceCount(&n_dev);
    if (e != cudaSuccess || n_dev == 0)
        return fail(nullptr, STOAT_GPU_ERR_CUDA,
                    std::string("stoat_gpu_create: no CUDA device (") + cudaGetErrorString(e) +
                        "); this library has no CPU fallback");
    if (device < 0 || device >= n_dev) return fail(nullptr, STOAT_GPU_ERR_ARG, "stoat_gpu_create: bad device index");
    cudaDeviceProp prop{};
    if ((e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess) return fail_cuda(nullptr, e, "cudaGetDeviceProperties");
    if (prop.major != 10)
        return fail(nullptr, STOAT_GPU_ERR_CUDA,
                    "stoat_gpu_create: device is sm_" + std::to_string(prop.major) + std::to_string(prop.minor) +
                        ", this library is built for sm_100a (B200) only");
    if ((e = cudaSetDevice(device)) != cudaSuccess) return fail_cuda(nullptr, e, "cudaSetDevice");
    stoat_gpu_ctx_full* full = new (std::nothrow) stoat_gpu_ctx_full();
    if (!full) return fail(nullptr, STOAT_GPU_ERR_OOM, "stoat_gpu_create: host allocation failed");
    ctx = full;
    ctx->device = device;
    ctx->num_sms = prop.multiProcessorCount;
    if ((e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking)) != cudaSuccess) {
        delete full;
        return fail_cuda(nullptr, e, "cudaStreamCreate");
    }
    for (auto& ev : ctx->ev)
        if ((e = cudaEventCreate(&ev)) != cudaSuccess) {
            delete full;
            return fail_cuda(nullptr, e, "cudaEventCreate");
        }
    // second stream + hand-over events of the overlapped binary run (stoat_gpu_binary_run_staged)
    if ((e = cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking)) != cudaSuccess) {
        delete full;
        return fail_cuda(nullptr, e, "cudaStreamCreate");
    }
    for (auto& ev : full->extra.chunk_done)
        if ((e = cudaEventCreateWithFlags(&ev, cudaEventDisableTiming)) != cudaSuccess) {
            delete full;
            return fail_cuda(nullptr, e, "cudaEventCreate");
        }
    for (auto& slot : full->extra.ring)
        for (auto& ev : slot)
            if ((e = cudaEventCreate(&ev)) != cudaSuccess) {
                delete full;
                return fail_cuda(nullptr, e, "cudaEventCreate");
            }
    const char* k1 = std::getenv("STOAT_GPU_K1");
    ctx->k1_variant = (k1 && std::strcmp(k1, "ldg") == 0) ? 1 : 0;
    const char* eq = std::getenv("STOAT_GPU_EQTL");
    // 0: tcgen05 int8 contraction (default); 2: the FP64 DMMA contraction (STOAT_GPU_EQTL=dmma); 1: general kernel only
    ctx->eqtl_variant = (eq && std::strcmp(eq, "general") == 0) ? 1 : (eq && std::strcmp(eq, "dmma") == 0) ? 2 : 0;
    *out = ctx;
    return STOAT_GPU_OK;
}

void stoat_gpu_destroy(stoat_gpu_ctx* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    if (ctx->copy_stream) cudaStreamSynchronize(ctx->copy_stream);
    DevBuf* bufs[] = {&ctx->rows, &ctx->raw_stage[0], &ctx->raw_stage[1], &ctx->case_mask, &ctx->is_case_dev, &ctx->y,
                      &ctx->covar, &ctx->d_snarl_path_off, &ctx->d_path_edge_off, &ctx->d_edge_row, &ctx->d_rec,
                      &ctx->d_blob, &ctx->d_big, &ctx->d_counters, &ctx->counts, &ctx->total_sum, &ctx->filtered, &ctx->filtered_bin,
                      &ctx->p_a, &ctx->p_b, &ctx->code_a, &ctx->code_b, &ctx->fisher_tab, &ctx->fisher_idx, &ctx->fisher_cnt, &ctx->q_allele, &ctx->q_nused, &ctx->q_p,
                      &ctx->q_r2, &ctx->q_beta, &ctx->q_se, &ctx->q_code, &ctx->e_expr, &ctx->e_gene_off, &ctx->e_gene_idx,
                      &ctx->e_min_p, &ctx->e_argmin, &ctx->e_nbelow, &ctx->e_yc, &ctx->e_tot, &ctx->e_tot2, &ctx->e_dcode,
                      &ctx->e_rec, &ctx->e_fb_list, &ctx->e_fb_count, &ctx->e_slot, &ctx->e_part_key, &ctx->e_part_gene,
                      &ctx->e_part_cnt};
    for (DevBuf* b : bufs) b->release();
    CtxExtra* x = extra_of(ctx);
    x->q.yc.release();
    x->q.ys.release();
    x->q.mom.release();
    x->q.part_i.release();
    x->q.part_d.release();
    x->q.slow.release();
    x->q.yslices.release();
    x->q.lean_rec.release();
    x->q.tail_t.release();
    x->q.tail_df.release();
    x->q.tail_x.release();
    x->bh.release();
    x->fmt.release();
    for (DevBuf* b : {&x->um_scale, &x->um_b, &x->um_a, &x->um_probe}) b->release();
    x->gather_idx.release();
    x->gather_out.release();
    for (int i = 0; i < kHostRing; ++i) {
        if (x->host_ring[i]) cudaFreeHost(x->host_ring[i]);
        if (x->host_ring_free[i]) cudaEventDestroy(x->host_ring_free[i]);
    }
    for (DevBuf* b : {&x->lg.counts, &x->lg.list, &x->lg.n_list, &x->lg.allele, &x->lg.filtered, &x->lg.n_used, &x->lg.p,
                      &x->lg.beta, &x->lg.se, &x->lg.code})
        b->release();
    for (auto& ev : ctx->ev)
        if (ev) cudaEventDestroy(ev);
    for (auto& ev : x->chunk_done)
        if (ev) cudaEventDestroy(ev);
    for (auto& slot : x->ring)
        for (auto& ev : slot)
            if (ev) cudaEventDestroy(ev);
    if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete static_cast<stoat_gpu_ctx_full*>(ctx);
}

// ---- matrix ----------------------------------------------------------------------------------------
static int set_matrix_common(stoat_gpu_ctx* ctx, const uint8_t* bits, bool on_device, uint64_t n_rows,
                             uint64_t n_cols) {
    if (!ctx) return STOAT_GPU_ERR_ARG;
    if (n_cols == 0 || (n_cols & 1)) return fail(ctx, STOAT_GPU_ERR_ARG, "set_matrix: n_cols must be even and > 0");
    if (n_cols >= (1ull << 32) || n_rows >= 0xFFFFFFFFull)
        return fail(ctx, STOAT_GPU_ERR_ARG, "set_matrix: dimensions exceed 32-bit row / column indices");
    if (n_rows && !bits) return fail(ctx, STOAT_GPU_ERR_ARG, "set_matrix: bits is NULL");
    if (on_device && (reinterpret_cast<uintptr_t>(bits) & 3))
        return fail(ctx, STOAT_GPU_ERR_ARG, "set_matrix_dev: device pointer must be 4-byte aligned");
    CK(cudaSetDevice(ctx->device));
    if (int rc = finalize_timing(ctx)) return rc;
    CtxExtra* x = extra_of(ctx);
    const uint64_t pitch = pitch_for(n_cols);
    const uint64_t raw_bytes = (n_rows * n_cols + 7) / 8;
    ctx->have_matrix = false;
    invalidate_derived(ctx);
    CK(ctx->rows.reserve(std::max<uint64_t>(n_rows * pitch, 16)));
    ctx->timing = stoat_gpu_timing{};
    uint32_t launches = 0;
    CK(cudaEventRecord(ctx->ev[0], ctx->stream));
    if (on_device) {
        CK(cudaEventRecord(ctx->ev[1], ctx->stream));
        CK(launch_repack_rows(bits, raw_bytes, 0, n_rows, n_cols, pitch, ctx->rows.as<uint8_t>(), ctx->stream));
        launches = n_rows ? 1 : 0;
    } else {
        // The reference-layout bytes go up in chunks of whole rows through two staging buffers: the copy of chunk k + 1
        // (copy stream) overlaps the repack of chunk k (main stream), and the staging never holds more than two chunks.
        // A chunk starts at a multiple of 128 rows, so its first bit sits on a 16-byte boundary of the bit stream and the
        // repack kernel can address the staging buffer as if it were the whole matrix (base pointer shifted back).
        // A pageable source (the reference's std::vector<uint8_t>, src/matrix.hpp:53) would make every cudaMemcpyAsync a
        // synchronous, single-threaded staged copy (measured 11 GB/s on the B200 box against 55 GB/s from pinned memory):
        // such a source goes through a ring of three pinned 64 MB buffers that a few host threads fill while the previous
        // buffer is on the wire.
        bool pageable = false;
        {
            cudaPointerAttributes at{};
            if (n_rows && cudaPointerGetAttributes(&at, bits) == cudaSuccess) pageable = at.type == cudaMemoryTypeUnregistered;
            (void)cudaGetLastError();
            if (const char* e = std::getenv("STOAT_GPU_UPLOAD_RING")) pageable = pageable && std::atoi(e) != 0;
        }
        uint64_t chunk_bytes = pageable ? kHostRingBytes : (256ull << 20);
        if (const char* e = std::getenv("STOAT_GPU_UPLOAD_CHUNK_MB")) chunk_bytes = std::max<uint64_t>(1, uint64_t(std::atoll(e))) << 20;
        if (pageable) {
            chunk_bytes = std::min<uint64_t>(chunk_bytes, kHostRingBytes);
            for (int i = 0; i < kHostRing; ++i) {
                if (!x->host_ring[i]) CK(cudaHostAlloc(&x->host_ring[i], kHostRingBytes + 64, cudaHostAllocDefault));
                if (!x->host_ring_free[i]) CK(cudaEventCreateWithFlags(&x->host_ring_free[i], cudaEventDisableTiming));
            }
        }
        uint64_t rows_per = std::max<uint64_t>(128, (chunk_bytes * 8 / n_cols) / 128 * 128);
        if (const char* e = std::getenv("STOAT_GPU_UPLOAD_CHUNK_ROWS"))  // tests: many chunks of a small matrix
            rows_per = std::max<uint64_t>(128, (uint64_t(std::atoll(e)) + 127) / 128 * 128);
        if (rows_per >= n_rows) rows_per = std::max<uint64_t>(n_rows, 1);
        const uint64_t n_chunks = n_rows ? (n_rows + rows_per - 1) / rows_per : 0;
        for (uint64_t k = 0; k < n_chunks; ++k) {
            const uint64_t r0 = k * rows_per, r1 = std::min(n_rows, r0 + rows_per);
            const uint64_t b0 = r0 * n_cols / 8;  // exact: r0 is a multiple of 128 (or 0)
            const uint64_t b1 = std::min(raw_bytes, (r1 * n_cols + 7) / 8);
            DevBuf& st = ctx->raw_stage[k & 1];
            if (k < 2) CK(st.reserve(std::min(raw_bytes, (rows_per * n_cols + 7) / 8) + 64));
            if (k >= 2) CK(cudaStreamWaitEvent(ctx->copy_stream, x->chunk_done[k & 1], 0));  // the buffer's previous repack
            else CK(cudaStreamWaitEvent(ctx->copy_stream, ctx->ev[0], 0));
            const uint8_t* src = bits + b0;
            if (pageable) {
                uint8_t* pin = static_cast<uint8_t*>(x->host_ring[k % kHostRing]);
                if (k >= uint64_t(kHostRing)) CK(cudaEventSynchronize(x->host_ring_free[k % kHostRing]));  // its last copy left
                parallel_memcpy(pin, src, b1 - b0);
                src = pin;
            }
            CK(cudaMemcpyAsync(st.p, src, b1 - b0, cudaMemcpyHostToDevice, ctx->copy_stream));
            if (pageable) CK(cudaEventRecord(x->host_ring_free[k % kHostRing], ctx->copy_stream));
            CK(cudaEventRecord(x->chunk_done[2 + (k & 1)], ctx->copy_stream));
            CK(cudaStreamWaitEvent(ctx->stream, x->chunk_done[2 + (k & 1)], 0));
            CK(launch_repack_rows(st.as<uint8_t>() - b0, raw_bytes, r0, r1, n_cols, pitch, ctx->rows.as<uint8_t>(),
                                  ctx->stream));
            CK(cudaEventRecord(x->chunk_done[k & 1], ctx->stream));
            ++launches;
        }
        CK(cudaEventRecord(ctx->ev[1], ctx->stream));
    }
    CK(cudaEventRecord(ctx->ev[2], ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    if (on_device) {
        CK(cudaEventElapsedTime(&ctx->timing.h2d_ms, ctx->ev[0], ctx->ev[1]));
        CK(cudaEventElapsedTime(&ctx->timing.repack_ms, ctx->ev[1], ctx->ev[2]));
    } else {
        CK(cudaEventElapsedTime(&ctx->timing.h2d_ms, ctx->ev[0], ctx->ev[2]));  // upload and repack overlap: one figure
        ctx->timing.repack_ms = 0.f;
    }
    ctx->timing.launches = launches;
    ctx->n_rows = n_rows;
    ctx->n_cols = n_cols;
    ctx->pitch = pitch;
    ctx->have_matrix = true;
    // phenotype masks depend on the pitch: they must be set again after a matrix of another width
    if (ctx->have_binary && ctx->n_samples_binary * 2 != n_cols) ctx->have_binary = false;
    if (ctx->have_quant && ctx->n_samples_quant * 2 != n_cols) ctx->have_quant = false;
    return STOAT_GPU_OK;
}

int stoat_gpu_set_matrix(stoat_gpu_ctx* ctx, const uint8_t* bits, uint64_t n_rows, uint64_t n_cols) {
    return set_matrix_common(ctx, bits, false, n_rows, n_cols);
}
int stoat_gpu_set_matrix_dev(stoat_gpu_ctx* ctx, const uint8_t* bits_dev, uint64_t n_rows, uint64_t n_cols) {
    return set_matrix_common(ctx, bits_dev, true, n_rows, n_cols);
}

int stoat_gpu_set_matrix_pairs(stoat_gpu_ctx* ctx, const uint32_t* rows, const uint32_t* cols, uint64_t n_pairs,
                               uint64_t n_rows, uint64_t n_cols) {
    if (!ctx) return STOAT_GPU_ERR_ARG;
    if (n_cols == 0 || (n_cols & 1)) return fail(ctx, STOAT_GPU_ERR_ARG, "set_matrix_pairs: n_cols must be even and > 0");
    if (n_cols >= (1ull << 32) || n_rows >= 0xFFFFFFFFull)
        return fail(ctx, STOAT_GPU_ERR_ARG, "set_matrix_pairs: dimensions exceed 32-bit row / column indices");
    if (n_pairs && (!rows || !cols)) return fail(ctx, STOAT_GPU_ERR_ARG, "set_matrix_pairs: rows / cols is NULL");
    CK(cudaSetDevice(ctx->device));
    if (int rc = finalize_timing(ctx)) return rc;
    const uint64_t pitch = pitch_for(n_cols);
    ctx->have_matrix = false;
    invalidate_derived(ctx);
    CK(ctx->rows.reserve(std::max<uint64_t>(n_rows * pitch, 16)));
    CK(ctx->d_counters.reserve(32));
    ctx->timing = stoat_gpu_timing{};
    CK(cudaEventRecord(ctx->ev[0], ctx->stream));
    CK(cudaMemsetAsync(ctx->rows.p, 0, std::max<uint64_t>(n_rows * pitch, 16), ctx->stream));
    CK(cudaMemsetAsync(ctx->d_counters.as<uint32_t>() + 1, 0, 4, ctx->stream));  // the scatter's error flag only
    extra_of(ctx)->q.lean_key = 0;
    // pairs go up in chunks through the two staging buffers: the copy of chunk k + 1 overlaps the scatter of chunk k
    const uint64_t chunk = 1ull << 25;  // 32M pairs = 256 MB
    CtxExtra* x = extra_of(ctx);
    float scatter_ms = 0.f;
    for (uint64_t p0 = 0, k = 0; p0 < n_pairs; p0 += chunk, ++k) {
        const uint64_t n = std::min(chunk, n_pairs - p0);
        DevBuf& st = ctx->raw_stage[k & 1];
        CK(st.reserve(std::min(chunk, n_pairs) * 8));
        uint32_t* d_rows = st.as<uint32_t>();
        uint32_t* d_cols = d_rows + n;
        if (k >= 2) CK(cudaStreamWaitEvent(ctx->copy_stream, x->chunk_done[k & 1], 0));  // the buffer's previous scatter
        CK(cudaMemcpyAsync(d_rows, rows + p0, n * 4, cudaMemcpyHostToDevice, ctx->copy_stream));
        CK(cudaMemcpyAsync(d_cols, cols + p0, n * 4, cudaMemcpyHostToDevice, ctx->copy_stream));
        CK(cudaEventRecord(x->chunk_done[2 + (k & 1)], ctx->copy_stream));
        CK(cudaStreamWaitEvent(ctx->stream, x->chunk_done[2 + (k & 1)], 0));
        CK(launch_scatter_bits(d_rows, d_cols, n, n_rows, n_cols, pitch, ctx->rows.as<uint8_t>(),
                               ctx->d_counters.as<uint32_t>() + 1, ctx->num_sms, ctx->stream));
        CK(cudaEventRecord(x->chunk_done[k & 1], ctx->stream));
    }
    CK(cudaEventRecord(ctx->ev[1], ctx->stream));
    CK(cudaMemcpyAsync(x->counters_host, ctx->d_counters.p, 16, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaEventElapsedTime(&scatter_ms, ctx->ev[0], ctx->ev[1]));
    ctx->timing.h2d_ms = scatter_ms;  // upload and scatter overlap: one figure
    ctx->timing.repack_ms = 0.f;
    ctx->timing.launches = uint32_t((n_pairs + chunk - 1) / chunk);
    ctx->have_matrix = false;
    if (x->counters_host[1])
        return fail(ctx, STOAT_GPU_ERR_ARG, "set_matrix_pairs: a (row, column) pair lies outside the matrix");
    ctx->n_rows = n_rows;
    ctx->n_cols = n_cols;
    ctx->pitch = pitch;
    ctx->have_matrix = true;
    if (ctx->have_binary && ctx->n_samples_binary * 2 != n_cols) ctx->have_binary = false;
    if (ctx->have_quant && ctx->n_samples_quant * 2 != n_cols) ctx->have_quant = false;
    return STOAT_GPU_OK;
}

// ---- post-processing ----------------------------------------------------------------------------------
int stoat_gpu_bh_adjust(stoat_gpu_ctx* ctx, const double* p, uint64_t n, double* p_adjusted) {
    if (!ctx) return STOAT_GPU_ERR_ARG;
    if (n == 0) return STOAT_GPU_OK;  // src/post_processing.cpp:9
    if (!p || !p_adjusted) return fail(ctx, STOAT_GPU_ERR_ARG, "bh_adjust: NULL argument");
    if (n >= (1ull << 31)) return fail(ctx, STOAT_GPU_ERR_ARG, "bh_adjust: more than 2^31 - 1 p-values");
    for (uint64_t i = 0; i < n; ++i)
        if (!(p[i] >= 0.0) || !std::isfinite(p[i]))
            return fail(ctx, STOAT_GPU_ERR_ARG, "bh_adjust: p-values must be finite and >= 0 (drop the NA rows first)");
    CK(cudaSetDevice(ctx->device));
    CtxExtra* x = extra_of(ctx);
    const size_t temp = bh_temp_bytes(n);
    // one scratch block: p | out | keys | a | b (8 n each) | idx_in | idx_sorted (4 n each) | temp
    const size_t need = size_t(n) * (5 * 8 + 2 * 4) + 256 + temp;
    CK(x->bh.reserve(need));
    uint8_t* base = x->bh.as<uint8_t>();
    double* d_p = reinterpret_cast<double*>(base);
    double* d_out = d_p + n;
    uint64_t* d_keys = reinterpret_cast<uint64_t*>(d_out + n);
    double* d_a = reinterpret_cast<double*>(d_keys + n);
    double* d_b = d_a + n;
    uint32_t* d_i0 = reinterpret_cast<uint32_t*>(d_b + n);
    uint32_t* d_i1 = d_i0 + n;
    void* d_temp = base + ((size_t(n) * 48 + 255) & ~size_t(255));
    CK(cudaMemcpyAsync(d_p, p, n * 8, cudaMemcpyHostToDevice, ctx->stream));
    CK(launch_bh_adjust(d_p, n, d_out, d_keys, d_i0, d_i1, d_a, d_b, d_temp, temp, ctx->stream));
    CK(cudaMemcpyAsync(p_adjusted, d_out, n * 8, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return STOAT_GPU_OK;
}

// No counterpart in the reference (see include/stoat_gpu.h): Benjamini-Hochberg with the scratch in a file-scope device
// array of kernels_post.cu.  Nothing here allocates: no DevBuf::reserve, no cudaMalloc / cudaFree, CUB is given its storage.
int stoat_gpu_bh_adjust_static(stoat_gpu_ctx* ctx, const double* p, uint64_t n, double* p_adjusted) {
    if (!ctx) return STOAT_GPU_ERR_ARG;
    if (n == 0) return STOAT_GPU_OK;  // src/post_processing.cpp:9
    if (!p || !p_adjusted) return fail(ctx, STOAT_GPU_ERR_ARG, "bh_adjust_static: NULL argument");
    if (n > bh_static_max_n())
        return fail(ctx, STOAT_GPU_ERR_ARG,
                    "bh_adjust_static: n = " + std::to_string(n) + " exceeds the static scratch (STOAT_GPU_BH_STATIC_MAX_N = " +
                        std::to_string(bh_static_max_n()) + " p-values); use stoat_gpu_bh_adjust");
    CK(cudaSetDevice(ctx->device));
    uint32_t bad = 0;
    CK(launch_bh_adjust_static(ctx->device, p, n, p_adjusted, &bad, ctx->stream));
    if (bad)
        return fail(ctx, STOAT_GPU_ERR_ARG, "bh_adjust_static: p-values must be finite and >= 0 (drop the NA rows first)");
    return STOAT_GPU_OK;
}

uint64_t stoat_gpu_row_pitch(const stoat_gpu_ctx* ctx) { return ctx ? ctx->pitch : 0; }
uint64_t stoat_gpu_stream(const stoat_gpu_ctx* ctx) { return ctx ? reinterpret_cast<uint64_t>(ctx->stream) : 0; }

// Rows of this context's (pitched, local-column) matrix gathered into a context-owned device buffer: what a rank of a
// sample-sharded run contributes to the re-assembly of the snarls it deferred.  The buffer stays valid until the next call.
int stoat_gpu_gather_rows_dev(stoat_gpu_ctx* ctx, const uint32_t* rows, uint64_t n, uint8_t** out_dev, uint64_t* pitch) {
    if (!ctx || !out_dev || !pitch || (n && !rows)) return fail(ctx, STOAT_GPU_ERR_ARG, "gather_rows_dev: NULL argument");
    if (!ctx->have_matrix) return fail(ctx, STOAT_GPU_ERR_STATE, "gather_rows_dev: no matrix");
    for (uint64_t i = 0; i < n; ++i)
        if (rows[i] >= ctx->n_rows) return fail(ctx, STOAT_GPU_ERR_ARG, "gather_rows_dev: row out of range");
    CK(cudaSetDevice(ctx->device));
    CtxExtra* x = extra_of(ctx);
    CK(x->gather_idx.reserve(std::max<uint64_t>(n * 4, 16)));
    CK(x->gather_out.reserve(std::max<uint64_t>(n * ctx->pitch, 16)));
    if (n) {
        CK(cudaMemcpyAsync(x->gather_idx.p, rows, n * 4, cudaMemcpyHostToDevice, ctx->stream));
        CK(launch_gather_rows(ctx->rows.as<uint8_t>(), ctx->pitch, x->gather_idx.as<uint32_t>(), n, x->gather_out.as<uint8_t>(),
                              ctx->stream));
    }
    CK(cudaStreamSynchronize(ctx->stream));
    *out_dev = x->gather_out.as<uint8_t>();
    *pitch = ctx->pitch;
    return STOAT_GPU_OK;
}

// The matrix from rows that are already in the device layout (src_pitch >= ceil(n_cols / 8) bytes per row, a multiple of 16,
// every bit beyond n_cols zero): no repack pass.
int stoat_gpu_set_matrix_pitched_dev(stoat_gpu_ctx* ctx, const uint8_t* rows_dev, uint64_t n_rows, uint64_t n_cols,
                                     uint64_t src_pitch) {
    if (!ctx) return STOAT_GPU_ERR_ARG;
    if (n_cols == 0 || (n_cols & 1)) return fail(ctx, STOAT_GPU_ERR_ARG, "set_matrix_pitched_dev: n_cols must be even and > 0");
    if (n_cols >= (1ull << 32) || n_rows >= 0xFFFFFFFFull)
        return fail(ctx, STOAT_GPU_ERR_ARG, "set_matrix_pitched_dev: dimensions exceed 32-bit row / column indices");
    if (n_rows && !rows_dev) return fail(ctx, STOAT_GPU_ERR_ARG, "set_matrix_pitched_dev: rows is NULL");
    if (src_pitch * 8 < n_cols || (src_pitch & 15)) return fail(ctx, STOAT_GPU_ERR_ARG, "set_matrix_pitched_dev: bad pitch");
    CK(cudaSetDevice(ctx->device));
    if (int rc = finalize_timing(ctx)) return rc;
    const uint64_t pitch = pitch_for(n_cols);
    ctx->have_matrix = false;
    invalidate_derived(ctx);
    CK(ctx->rows.reserve(std::max<uint64_t>(n_rows * pitch, 16)));
    ctx->timing = stoat_gpu_timing{};
    if (n_rows) {
        CK(cudaMemsetAsync(ctx->rows.p, 0, n_rows * pitch, ctx->stream));
        CK(cudaMemcpy2DAsync(ctx->rows.p, pitch, rows_dev, src_pitch, std::min(pitch, src_pitch), n_rows,
                             cudaMemcpyDeviceToDevice, ctx->stream));
    }
    CK(cudaStreamSynchronize(ctx->stream));
    ctx->n_rows = n_rows;
    ctx->n_cols = n_cols;
    ctx->pitch = pitch;
    ctx->have_matrix = true;
    if (ctx->have_binary && ctx->n_samples_binary * 2 != n_cols) ctx->have_binary = false;
    if (ctx->have_quant && ctx->n_samples_quant * 2 != n_cols) ctx->have_quant = false;
    return STOAT_GPU_OK;
}

int stoat_gpu_read_row(stoat_gpu_ctx* ctx, uint64_t row, uint8_t* out, uint64_t out_bytes) {
    if (!ctx || !out) return STOAT_GPU_ERR_ARG;
    if (!ctx->have_matrix) return fail(ctx, STOAT_GPU_ERR_STATE, "read_row: no matrix");
    if (row >= ctx->n_rows || out_bytes < ctx->pitch) return fail(ctx, STOAT_GPU_ERR_ARG, "read_row: bad row / size");
    CK(cudaSetDevice(ctx->device));
    CK(cudaMemcpyAsync(out, ctx->rows.as<uint8_t>() + row * ctx->pitch, ctx->pitch, cudaMemcpyDeviceToHost,
                       ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return STOAT_GPU_OK;
}

// ---- phenotypes ------------------------------------------------------------------------------------
int stoat_gpu_set_binary_phenotype(stoat_gpu_ctx* ctx, const uint8_t* is_case, uint64_t n_samples) {
    if (!ctx || !is_case) return fail(ctx, STOAT_GPU_ERR_ARG, "set_binary_phenotype: NULL argument");
    if (!ctx->have_matrix) return fail(ctx, STOAT_GPU_ERR_STATE, "set_binary_phenotype: set the matrix first");
    if (n_samples * 2 != ctx->n_cols)
        return fail(ctx, STOAT_GPU_ERR_ARG, "set_binary_phenotype: n_samples must be n_cols / 2");
    CK(cudaSetDevice(ctx->device));
    CK(ctx->is_case_dev.reserve(n_samples));
    CK(ctx->case_mask.reserve(ctx->pitch));
    CK(cudaMemcpyAsync(ctx->is_case_dev.p, is_case, n_samples, cudaMemcpyHostToDevice, ctx->stream));
    CK(launch_build_case_mask(ctx->is_case_dev.as<uint8_t>(), n_samples, ctx->pitch, ctx->case_mask.as<uint8_t>(),
                              ctx->stream));
    uint64_t cases = 0;
    for (uint64_t i = 0; i < n_samples; ++i) cases += is_case[i] ? 1 : 0;
    CK(cudaStreamSynchronize(ctx->stream));
    ctx->n_case_hap = uint32_t(cases * 2);
    ctx->n_samples_binary = n_samples;
    ctx->have_binary = true;
    return STOAT_GPU_OK;
}

// shared by the whole-phenotype and the sample-shard entry points: `mean`, `s1`, `s2` (sums of y - mean and its square)
// and `vmax` (max |y - mean|) describe the WHOLE phenotype, y[0 .. n_samples) is what this GPU holds
static int set_quant_common(stoat_gpu_ctx* ctx, const double* y, uint64_t n_samples, double mean, double s1, double s2,
                            double vmax) {
    CK(cudaSetDevice(ctx->device));
    CtxExtra* x = extra_of(ctx);
    x->q.sy_all = s1;
    x->q.syy_all = s2;
    const uint32_t n_words = uint32_t(ctx->pitch / 4);
    CK(ctx->y.reserve(n_samples * 8));
    CK(x->q.yc.reserve(n_samples * 8));
    CK(x->q.ys.reserve(size_t(16) * n_words * 8));
    CK(cudaMemcpyAsync(ctx->y.p, y, n_samples * 8, cudaMemcpyHostToDevice, ctx->stream));
    CK(launch_centre_permute(ctx->y.as<double>(), uint32_t(n_samples), mean, n_words, x->q.yc.as<double>(),
                             x->q.ys.as<double>(), ctx->stream));
    // K3M's operand: the centred phenotype as 8 signed 7-bit slices of one fixed-point scale,
    //     y - mean = scale * sum_t q_t * 128^-(t+1)   (every step below is exact in FP64; residual <= 2^-57 scale),
    // laid out as the B fragments of its integer MMAs (csrc/quant_mma.cuh): [round c][word j][mma m][lane (g, t)][8 B];
    // lane (g, t) works on the 16-byte chunk 4 c + t of a row, i.e. on word (4 c + t) * 4 + j, and register b, byte bb
    // of its fragment is slice g of that word's sample 4 bb + 2 m + b.
    {
        int ex = 0;
        if (vmax > 0.0 && std::isfinite(vmax)) std::frexp(vmax, &ex);  // vmax = f * 2^ex, f in [0.5, 1)
        x->q.y_scale = std::ldexp(1.0, ex + 1);                        // |y - mean| / scale < 0.5
        std::vector<int8_t> tab(quant_yslices_bytes(ctx->pitch), 0);
        for (uint64_t i = 0; i < n_samples; ++i) {
            double r = std::ldexp(y[i] - mean, -(ex + 1));
            int8_t qs[8];
            for (int t = 0; t < 8; ++t) {
                r *= 128.0;
                const double qv = std::nearbyint(r);
                qs[t] = (std::isfinite(qv) && std::fabs(qv) <= 127.0) ? int8_t(qv) : int8_t(0);
                r -= qv;
            }
            const uint64_t word = i / 16, k = i % 16;  // k = 4 bb + 2 m + b
            const uint64_t chunk = word / 4, j = word % 4, c = chunk / 4, t = chunk % 4;
            const uint64_t bb = k / 4, m = (k % 4) / 2, b = k % 2;
            for (uint64_t g = 0; g < 8; ++g)
                tab[((((c * 4 + j) * 2 + m) * 32 + (g * 4 + t)) * 8) + b * 4 + bb] = qs[g];
        }
        CK(x->q.yslices.reserve(std::max<size_t>(tab.size(), 16)));
        CK(cudaMemcpyAsync(x->q.yslices.p, tab.data(), tab.size(), cudaMemcpyHostToDevice, ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));
    }
    CK(cudaStreamSynchronize(ctx->stream));
    ctx->n_samples_quant = n_samples;
    ctx->have_quant = true;
    ctx->have_quant_result = false;
    return STOAT_GPU_OK;
}

int stoat_gpu_set_quantitative_phenotype(stoat_gpu_ctx* ctx, const double* y, uint64_t n_samples) {
    if (!ctx || !y) return fail(ctx, STOAT_GPU_ERR_ARG, "set_quantitative_phenotype: NULL argument");
    if (!ctx->have_matrix) return fail(ctx, STOAT_GPU_ERR_STATE, "set_quantitative_phenotype: set the matrix first");
    if (n_samples * 2 != ctx->n_cols)
        return fail(ctx, STOAT_GPU_ERR_ARG, "set_quantitative_phenotype: n_samples must be n_cols / 2");
    // centring constant and the all-sample sums of the centred values (O(n), once per phenotype)
    long double acc = 0.0L;
    for (uint64_t i = 0; i < n_samples; ++i) acc += y[i];
    const double mean = double(acc / (long double)n_samples);
    long double s1 = 0.0L, s2 = 0.0L;
    double vmax = 0.0;
    for (uint64_t i = 0; i < n_samples; ++i) {
        const double v = y[i] - mean;
        s1 += v;
        s2 += (long double)v * v;
        vmax = std::max(vmax, std::fabs(v));
    }
    return set_quant_common(ctx, y, n_samples, mean, double(s1), double(s2), vmax);
}

int stoat_gpu_set_quantitative_phenotype_shard(stoat_gpu_ctx* ctx, const double* y_local, uint64_t n_local, double mean,
                                               double sum_centred, double sum_centred_sq, double max_abs_centred) {
    if (!ctx) return STOAT_GPU_ERR_ARG;
    if (!y_local) return fail(ctx, STOAT_GPU_ERR_ARG, "set_quantitative_phenotype_shard: y is NULL");
    if (!ctx->have_matrix) return fail(ctx, STOAT_GPU_ERR_STATE, "set_quantitative_phenotype_shard: set the matrix first");
    if (n_local * 2 != ctx->n_cols)
        return fail(ctx, STOAT_GPU_ERR_ARG, "set_quantitative_phenotype_shard: n_local must be n_cols / 2");
    return set_quant_common(ctx, y_local, n_local, mean, sum_centred, sum_centred_sq, max_abs_centred);
}

int stoat_gpu_set_covariates(stoat_gpu_ctx* ctx, const double* covar, uint64_t n_samples, uint64_t n_covar) {
    if (!ctx) return STOAT_GPU_ERR_ARG;
    if (n_covar == 0) {
        ctx->n_covar = 0;
        ctx->n_samples_covar = 0;
        return STOAT_GPU_OK;
    }
    if (!covar) return fail(ctx, STOAT_GPU_ERR_ARG, "set_covariates: covar is NULL");
    if (n_covar + 2 > STOAT_GPU_MAX_OLS_FEATURES)
        return fail(ctx, STOAT_GPU_ERR_ARG, "set_covariates: too many covariates for STOAT_GPU_MAX_OLS_FEATURES");
    CK(cudaSetDevice(ctx->device));
    CK(ctx->covar.reserve(n_samples * n_covar * 8));
    CK(cudaMemcpyAsync(ctx->covar.p, covar, n_samples * n_covar * 8, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    if (quant_covar_fast(uint32_t(n_covar))) {
        // The reference pairs covariate row r with the r-th row of the filtered table (stats_test.cpp:450-452), so 1'C and
        // C'C of a snarl are sums over the FIRST n_used covariate rows: prefix[r] = those sums over rows [0, r), running
        // down the rows as the reference's own accumulation does.  Layout per r: sum c_j (Q), then sum c_j c_k, j <= k.
        const uint64_t Q = n_covar, per = Q + Q * (Q + 1) / 2;
        std::vector<double> pre((n_samples + 1) * per + n_samples * Q, 0.0);  // + the covariates by column, for K3C
        for (uint64_t r = 0; r < n_samples; ++r) {
            const double* row = covar + r * Q;
            const double* in = pre.data() + r * per;
            double* o = pre.data() + (r + 1) * per;
            uint64_t e = Q;
            for (uint64_t j = 0; j < Q; ++j) {
                o[j] = in[j] + row[j];
                for (uint64_t k = j; k < Q; ++k, ++e) o[e] = in[e] + row[j] * row[k];
            }
        }
        double* by_col = pre.data() + (n_samples + 1) * per;
        for (uint64_t r = 0; r < n_samples; ++r)
            for (uint64_t j = 0; j < Q; ++j) by_col[j * n_samples + r] = covar[r * Q + j];
        CtxExtra* x = extra_of(ctx);
        CK(x->q.cov_prefix.reserve(pre.size() * 8));
        CK(cudaMemcpy(x->q.cov_prefix.p, pre.data(), pre.size() * 8, cudaMemcpyHostToDevice));
    }
    ctx->n_covar = n_covar;
    ctx->n_samples_covar = n_samples;
    return STOAT_GPU_OK;
}

// ---- staging of a batch ----------------------------------------------------------------------------
int stoat_gpu_stage_paths(stoat_gpu_ctx* ctx, const stoat_gpu_paths* paths) {
    if (!ctx || !paths) return fail(ctx, STOAT_GPU_ERR_ARG, "stage_paths: NULL argument");
    if (!ctx->have_matrix) return fail(ctx, STOAT_GPU_ERR_STATE, "stage_paths: set the matrix first");
    const uint64_t S = paths->n_snarls;
    if (S && (!paths->snarl_path_off || !paths->path_edge_off))
        return fail(ctx, STOAT_GPU_ERR_ARG, "stage_paths: offset arrays are NULL");
    if (S >= 0xFFFFFFFFull) return fail(ctx, STOAT_GPU_ERR_ARG, "stage_paths: too many snarls for one batch");
    const uint64_t P = S ? paths->snarl_path_off[S] : 0;
    if (S && paths->snarl_path_off[0] != 0) return fail(ctx, STOAT_GPU_ERR_ARG, "stage_paths: snarl_path_off[0] != 0");
    const uint64_t E = P ? paths->path_edge_off[P] : 0;
    if (P && paths->path_edge_off[0] != 0) return fail(ctx, STOAT_GPU_ERR_ARG, "stage_paths: path_edge_off[0] != 0");
    if (P >= 0xFFFFFFFFull || E >= 0xFFFFFFFFull)
        return fail(ctx, STOAT_GPU_ERR_ARG, "stage_paths: more than 2^32 paths / edges in one batch");
    if (E && !paths->edge_row) return fail(ctx, STOAT_GPU_ERR_ARG, "stage_paths: edge_row is NULL");
    for (uint64_t s = 0; s < S; ++s)
        if (paths->snarl_path_off[s + 1] < paths->snarl_path_off[s])
            return fail(ctx, STOAT_GPU_ERR_ARG, "stage_paths: snarl_path_off is not monotone");
    for (uint64_t p = 0; p < P; ++p)
        if (paths->path_edge_off[p + 1] < paths->path_edge_off[p])
            return fail(ctx, STOAT_GPU_ERR_ARG, "stage_paths: path_edge_off is not monotone");
    CK(cudaSetDevice(ctx->device));
    if (int rc = finalize_timing(ctx)) return rc;
    CtxExtra* x = extra_of(ctx);
    ctx->have_batch = false;
    ctx->have_binary_counts = ctx->have_binary_result = ctx->have_quant_result = ctx->have_eqtl_summary = false;
    extra_of(ctx)->q.have_partials = false;
    extra_of(ctx)->lg.have_result = false;
    ++x->stage_gen;  // lean records of the previous staging are stale
    CK(ctx->d_snarl_path_off.reserve((S + 1) * 8));
    CK(ctx->d_path_edge_off.reserve((P + 1) * 8));
    CK(ctx->d_edge_row.reserve(std::max<uint64_t>(E * 4, 16)));
    CK(ctx->d_rec.reserve(std::max<uint64_t>(S * sizeof(SnarlRec), 32)));
    const uint64_t blob_bytes = 16 * (2 * S + ((P + E) >> 4) + 2 + kMaxBlob16);
    CK(ctx->d_blob.reserve(blob_bytes));
    CK(ctx->d_big.reserve(std::max<uint64_t>(S * 4, 16)));
    CK(ctx->d_counters.reserve(32));  // [0..3] see BatchDev, [4] thirds appended by K3M, [5] [6] three-path / thirds records, [7] K4b work counter
    CK(cudaEventRecord(ctx->ev[0], ctx->stream));
    const uint64_t zero = 0;
    if (S) {
        CK(cudaMemcpyAsync(ctx->d_snarl_path_off.p, paths->snarl_path_off, (S + 1) * 8, cudaMemcpyHostToDevice, ctx->stream));
        CK(cudaMemcpyAsync(ctx->d_path_edge_off.p, P ? paths->path_edge_off : &zero, (P + 1) * 8, cudaMemcpyHostToDevice,
                           ctx->stream));
        if (E) CK(cudaMemcpyAsync(ctx->d_edge_row.p, paths->edge_row, E * 4, cudaMemcpyHostToDevice, ctx->stream));
    }
    CK(cudaMemsetAsync(ctx->d_counters.p, 0, 32, ctx->stream));
    CK(cudaEventRecord(ctx->ev[1], ctx->stream));
    BatchDev& b = ctx->batch;
    b.n_snarls = S;
    b.n_paths = P;
    b.n_edges = E;
    b.snarl_path_off = ctx->d_snarl_path_off.as<uint64_t>();
    b.path_edge_off = ctx->d_path_edge_off.as<uint64_t>();
    b.edge_row = ctx->d_edge_row.as<uint32_t>();
    b.rec = ctx->d_rec.as<SnarlRec>();
    b.blob = ctx->d_blob.as<uint8_t>();
    b.big_list = ctx->d_big.as<uint32_t>();
    b.counters = ctx->d_counters.as<uint32_t>();
    CK(launch_binary_prep(b, ctx->n_rows, ctx->stream));
    CK(cudaMemcpyAsync(x->counters_host, ctx->d_counters.p, 16, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[1]));
    ctx->timing = stoat_gpu_timing{};
    ctx->timing.h2d_ms = ms;
    ctx->timing.launches = S ? 1 : 0;
    if (x->counters_host[1])
        return fail(ctx, STOAT_GPU_ERR_ARG, "stage_paths: edge_row holds a row index >= n_rows (and != MISSING_EDGE)");
    ctx->n_big = x->counters_host[0];
    x->max_stage_rows = x->counters_host[2];
    ctx->have_batch = true;
    return STOAT_GPU_OK;
}

// ---- binary ----------------------------------------------------------------------------------------
static int binary_reserve(stoat_gpu_ctx* ctx) {
    const BatchDev& b = ctx->batch;
    CK(ctx->counts.reserve(std::max<uint64_t>(2 * b.n_paths * 4, 16)));
    CK(ctx->total_sum.reserve(std::max<uint64_t>(b.n_snarls * 8, 16)));
    CK(ctx->filtered_bin.reserve(std::max<uint64_t>(b.n_snarls, 16)));
    CK(ctx->p_a.reserve(std::max<uint64_t>(b.n_snarls * 8, 16)));
    CK(ctx->p_b.reserve(std::max<uint64_t>(b.n_snarls * 8, 16)));
    CK(ctx->code_a.reserve(std::max<uint64_t>(b.n_snarls, 16)));
    CK(ctx->code_b.reserve(std::max<uint64_t>(b.n_snarls, 16)));
    CK(ctx->fisher_tab.reserve(std::max<uint64_t>(b.n_snarls * 16, 16)));
    CK(ctx->fisher_idx.reserve(std::max<uint64_t>(b.n_snarls * 4, 16)));
    CK(ctx->fisher_cnt.reserve(8 * kMaxRunChunks));  // a pair of work-list counters per chunk of an overlapped run
    return STOAT_GPU_OK;
}

int stoat_gpu_binary_counts_staged(stoat_gpu_ctx* ctx) {
    if (!ctx) return STOAT_GPU_ERR_ARG;
    if (!ctx->have_batch) return fail(ctx, STOAT_GPU_ERR_STATE, "binary: no staged batch");
    if (!ctx->have_binary) return fail(ctx, STOAT_GPU_ERR_STATE, "binary: no binary phenotype");
    CK(cudaSetDevice(ctx->device));
    int rc = binary_reserve(ctx);
    if (rc) return rc;
    CtxExtra* x = extra_of(ctx);
    ctx->timing.launches = 0;
    CK(cudaEventRecord(ctx->ev[2], ctx->stream));
    CK(launch_path_and_popc(matrix_view(ctx), ctx->case_mask.as<uint8_t>(), ctx->n_case_hap, ctx->batch, ctx->n_big,
                            x->max_stage_rows, ctx->counts.as<uint32_t>(), ctx->num_sms, ctx->k1_variant, ctx->stream,
                            &ctx->timing.launches));
    CK(cudaEventRecord(ctx->ev[3], ctx->stream));
    ctx->have_binary_counts = true;
    ctx->have_binary_result = false;
    return STOAT_GPU_OK;
}

int stoat_gpu_binary_counts_dev(stoat_gpu_ctx* ctx, uint32_t** counts_dev, uint64_t* n_words) {
    if (!ctx || !counts_dev || !n_words) return STOAT_GPU_ERR_ARG;
    if (!ctx->have_binary_counts) return fail(ctx, STOAT_GPU_ERR_STATE, "binary_counts_dev: run the count pass first");
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));  // the caller's collective runs on its own stream
    *counts_dev = ctx->counts.as<uint32_t>();
    *n_words = 2 * ctx->batch.n_paths;
    return STOAT_GPU_OK;
}

int stoat_gpu_binary_stats_staged(stoat_gpu_ctx* ctx, const stoat_gpu_thresholds* thr) {
    if (!ctx || !thr) return fail(ctx, STOAT_GPU_ERR_ARG, "binary_stats: NULL argument");
    if (!ctx->have_binary_counts) return fail(ctx, STOAT_GPU_ERR_STATE, "binary_stats: run the count pass first");
    CK(cudaSetDevice(ctx->device));
    CK(cudaEventRecord(ctx->ev[4], ctx->stream));
    CK(launch_binary_stats(ctx->batch, ctx->counts.as<uint32_t>(), *thr, ctx->total_sum.as<uint64_t>(),
                           ctx->filtered_bin.as<uint8_t>(), ctx->p_a.as<double>(), ctx->p_b.as<double>(),
                           ctx->code_a.as<uint8_t>(), ctx->code_b.as<uint8_t>(), ctx->fisher_tab.as<uint4>(),
                           ctx->fisher_idx.as<uint32_t>(), ctx->fisher_cnt.as<uint32_t>(), ctx->num_sms, ctx->stream));
    CK(cudaEventRecord(ctx->ev[5], ctx->stream));
    if (ctx->batch.n_snarls) ctx->timing.launches += 2;
    ctx->have_binary_result = true;
    return STOAT_GPU_OK;
}

int stoat_gpu_binary_run_staged(stoat_gpu_ctx* ctx, const stoat_gpu_thresholds* thr) {
    if (!ctx || !thr) return fail(ctx, STOAT_GPU_ERR_ARG, "binary: NULL argument");
    if (!ctx->have_batch) return fail(ctx, STOAT_GPU_ERR_STATE, "binary: no staged batch");
    if (!ctx->have_binary) return fail(ctx, STOAT_GPU_ERR_STATE, "binary: no binary phenotype");
    CK(cudaSetDevice(ctx->device));
    int rc = binary_reserve(ctx);
    if (rc) return rc;
    CtxExtra* x = extra_of(ctx);
    const uint64_t S = ctx->batch.n_snarls;
    // The count pass (K1) is HBM bound and leaves the FP64 pipe idle, the Fisher work list (K2b) is the opposite, so the
    // run CAN be cut into chunks of snarls with K1 of chunk c + 1 (main stream) overlapping K2a / K2b of chunk c (second
    // stream): STOAT_GPU_OVERLAP=<chunks>.  For the two to share an SM the count pass then runs in its 64-register build
    // (416 threads x 64 registers + four 128-thread Fisher blocks fit the register file).  See kDefaultRunChunks for
    // what that measures.
    int n_chunks = 1;
    if (S >= 65536) {
        n_chunks = kDefaultRunChunks;
        if (const char* e = std::getenv("STOAT_GPU_OVERLAP")) {
            const int v = std::atoi(e);
            n_chunks = v <= 1 ? 1 : (v > kMaxRunChunks ? kMaxRunChunks : v);
        }
    }
    bool lean = n_chunks > 1;
    if (const char* e = std::getenv("STOAT_GPU_K1_LEAN")) lean = std::atoi(e) != 0;
    if (int frc = finalize_timing(ctx)) return frc;
    ctx->timing.launches = 0;
    cudaEvent_t* ev = x->ring[x->run_seq % kRunRing];
    x->ring_kind[x->run_seq % kRunRing] = 1;
    const uint64_t per = (S + uint64_t(n_chunks) - 1) / uint64_t(n_chunks);
    cudaStream_t main_st = ctx->stream, side_st = n_chunks > 1 ? ctx->copy_stream : ctx->stream;
    CK(cudaEventRecord(ev[0], main_st));
    if (!S) {
        CK(cudaEventRecord(ev[1], main_st));
        CK(cudaEventRecord(ev[2], main_st));
    }
    for (int c = 0; c < n_chunks && S; ++c) {
        const uint64_t s0 = per * uint64_t(c);
        if (s0 >= S) break;
        const uint64_t cnt = std::min<uint64_t>(per, S - s0);
        CK(launch_path_and_popc(matrix_view(ctx), ctx->case_mask.as<uint8_t>(), ctx->n_case_hap, ctx->batch, ctx->n_big,
                                x->max_stage_rows, ctx->counts.as<uint32_t>(), ctx->num_sms, ctx->k1_variant, main_st,
                                &ctx->timing.launches, s0, cnt, lean));
        if (n_chunks > 1) {
            CK(cudaEventRecord(x->chunk_done[c], main_st));
            CK(cudaStreamWaitEvent(side_st, x->chunk_done[c], 0));
        } else {
            CK(cudaEventRecord(ev[1], main_st));
        }
        if (c == 0) CK(cudaEventRecord(ev[2], side_st));
        CK(launch_binary_stats(ctx->batch, ctx->counts.as<uint32_t>(), *thr, ctx->total_sum.as<uint64_t>(),
                               ctx->filtered_bin.as<uint8_t>(), ctx->p_a.as<double>(), ctx->p_b.as<double>(),
                               ctx->code_a.as<uint8_t>(), ctx->code_b.as<uint8_t>(), ctx->fisher_tab.as<uint4>(),
                               ctx->fisher_idx.as<uint32_t>(), ctx->fisher_cnt.as<uint32_t>() + 2 * c, ctx->num_sms,
                               side_st, s0, cnt, n_chunks > 1));
        ctx->timing.launches += 2;
    }
    if (n_chunks > 1 && S) CK(cudaEventRecord(ev[1], main_st));  // end of the last count pass
    CK(cudaEventRecord(ev[3], side_st));
    if (n_chunks > 1) CK(cudaStreamWaitEvent(main_st, ev[3], 0));
    ctx->have_binary_counts = true;
    ctx->have_binary_result = true;
    // nothing is read back here: the run is asynchronous, stoat_gpu_binary_fetch / stoat_gpu_last_timing wait for it
    ++x->run_seq;
    x->pending = 1;
    x->pending_chunks = n_chunks;
    return STOAT_GPU_OK;
}

int stoat_gpu_binary_fetch(stoat_gpu_ctx* ctx, stoat_gpu_binary_out* out) {
    if (!ctx || !out) return fail(ctx, STOAT_GPU_ERR_ARG, "binary_fetch: NULL argument");
    if (!ctx->have_binary_result) return fail(ctx, STOAT_GPU_ERR_STATE, "binary_fetch: nothing to fetch");
    CK(cudaSetDevice(ctx->device));
    const uint64_t S = ctx->batch.n_snarls, P = ctx->batch.n_paths;
    CK(cudaEventRecord(ctx->ev[6], ctx->stream));
    const cudaMemcpyKind k = cudaMemcpyDeviceToHost;
    if (P && out->g0) CK(cudaMemcpyAsync(out->g0, ctx->counts.as<uint32_t>(), P * 4, k, ctx->stream));
    if (P && out->g1) CK(cudaMemcpyAsync(out->g1, ctx->counts.as<uint32_t>() + P, P * 4, k, ctx->stream));
    if (S && out->total_sum) CK(cudaMemcpyAsync(out->total_sum, ctx->total_sum.p, S * 8, k, ctx->stream));
    if (S && out->filtered) CK(cudaMemcpyAsync(out->filtered, ctx->filtered_bin.p, S, k, ctx->stream));
    if (S && out->p_chi2) CK(cudaMemcpyAsync(out->p_chi2, ctx->p_a.p, S * 8, k, ctx->stream));
    if (S && out->p_fisher) CK(cudaMemcpyAsync(out->p_fisher, ctx->p_b.p, S * 8, k, ctx->stream));
    if (S && out->chi2_code) CK(cudaMemcpyAsync(out->chi2_code, ctx->code_a.p, S, k, ctx->stream));
    if (S && out->fisher_code) CK(cudaMemcpyAsync(out->fisher_code, ctx->code_b.p, S, k, ctx->stream));
    CK(cudaEventRecord(ctx->ev[7], ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    if (int rc = finalize_timing(ctx)) return rc;
    CK(cudaEventElapsedTime(&ctx->timing.d2h_ms, ctx->ev[6], ctx->ev[7]));
    return STOAT_GPU_OK;
}

int stoat_gpu_binary_batch(stoat_gpu_ctx* ctx, const stoat_gpu_paths* paths, const stoat_gpu_thresholds* thr,
                           stoat_gpu_binary_out* out) {
    if (!ctx || !paths || !thr || !out) return fail(ctx, STOAT_GPU_ERR_ARG, "binary_batch: NULL argument");
    int rc = stoat_gpu_stage_paths(ctx, paths);
    if (rc) return rc;
    const float h2d = ctx->timing.h2d_ms;
    rc = stoat_gpu_binary_run_staged(ctx, thr);
    if (rc) return rc;
    rc = stoat_gpu_binary_fetch(ctx, out);
    ctx->timing.h2d_ms = h2d;
    if (ctx->batch.n_snarls) ctx->timing.launches += 1;  // prep
    return rc;
}

// ---- quantitative ----------------------------------------------------------------------------------
static int quant_reserve(stoat_gpu_ctx* ctx) {
    const BatchDev& b = ctx->batch;
    CtxExtra* x = extra_of(ctx);
    CK(ctx->q_allele.reserve(std::max<uint64_t>(b.n_paths * 4, 16)));
    CK(ctx->filtered.reserve(std::max<uint64_t>(b.n_snarls, 16)));
    CK(ctx->q_nused.reserve(std::max<uint64_t>(b.n_snarls * 4, 16)));
    CK(ctx->q_p.reserve(std::max<uint64_t>(b.n_snarls * 8, 16)));
    CK(ctx->q_r2.reserve(std::max<uint64_t>(b.n_snarls * 8, 16)));
    CK(ctx->q_beta.reserve(std::max<uint64_t>(b.n_snarls * 8, 16)));
    CK(ctx->q_se.reserve(std::max<uint64_t>(b.n_snarls * 8, 16)));
    CK(ctx->q_code.reserve(std::max<uint64_t>(b.n_snarls, 16)));
    CK(x->q.mom.reserve(std::max<size_t>(quant_moments_bytes(b.n_snarls), 16)));
    const uint64_t slices = std::max<uint32_t>(quant_lean_slices(ctx->pitch, ctx->n_cols, x->max_stage_rows), 1);
    CK(x->q.part_i.reserve(std::max<uint64_t>(slices * 7 * b.n_snarls * 4, 16)));
    CK(x->q.part_d.reserve(std::max<uint64_t>(slices * 3 * b.n_snarls * 8, 16)));
    CK(x->q.slow.reserve(std::max<uint64_t>(b.n_snarls * 4, 16)));
    CK(x->q.lean_rec.reserve(std::max<uint64_t>(b.n_snarls * 80, 16)));  // [S] two-path records, [S][2] three-path, [S][2] thirds
    CK(x->q.tail_t.reserve(std::max<uint64_t>(b.n_snarls * 8, 16)));
    CK(x->q.tail_df.reserve(std::max<uint64_t>(b.n_snarls * 4, 16)));
    CK(x->q.tail_x.reserve(std::max<uint64_t>(b.n_snarls * 16, 16)));  // [2][S]: argument, starting state
    if (ctx->n_covar && quant_covar_fast(uint32_t(ctx->n_covar)))
        CK(x->q.cov_mom.reserve(std::max<size_t>(quant_covar_moments_bytes(b.n_snarls), 16)));
    return STOAT_GPU_OK;
}

// phase 0: the whole step; 1: moments only (K3M partial sums stay on the device); 2: solve on the (all-reduced) partials
static int quant_run(stoat_gpu_ctx* ctx, const stoat_gpu_thresholds* thr, int phase) {
    if (!ctx || !thr) return fail(ctx, STOAT_GPU_ERR_ARG, "quantitative: NULL argument");
    if (!ctx->have_batch) return fail(ctx, STOAT_GPU_ERR_STATE, "quantitative: no staged batch");
    if (!ctx->have_quant) return fail(ctx, STOAT_GPU_ERR_STATE, "quantitative: no quantitative phenotype");
    if (ctx->n_covar && ctx->n_samples_covar * 2 < ctx->n_cols)
        return fail(ctx, STOAT_GPU_ERR_ARG, "quantitative: fewer covariate rows than samples");
    CK(cudaSetDevice(ctx->device));
    if (phase != 0 && ctx->n_covar) return fail(ctx, STOAT_GPU_ERR_STATE, "quantitative: the sample-sharded phases do not take covariates");
    int rc = quant_reserve(ctx);
    if (rc) return rc;
    CtxExtra* x = extra_of(ctx);
    const uint64_t S = ctx->batch.n_snarls;
    if (phase == 2 && !x->q.have_partials) return fail(ctx, STOAT_GPU_ERR_STATE, "quantitative_solve_staged: run quantitative_moments_staged first");
    // the general-path list grows during the run: reset its length to what prep put there
    x->counters_host[0] = ctx->n_big;
    if (ctx->n_big == 0)
        CK(cudaMemsetAsync(ctx->d_counters.p, 0, 4, ctx->stream));
    else
        CK(cudaMemcpyAsync(ctx->d_counters.p, x->counters_host, 4, cudaMemcpyHostToDevice, ctx->stream));
    // filtered snarls keep NaN / NA outputs (the moments path writes them in K4, which visits every snarl)
    if (S && phase != 1 && !quant_fast_path(ctx->pitch, ctx->n_cols, x->max_stage_rows, ctx->k1_variant, uint32_t(ctx->n_covar))) {
        CK(cudaMemsetAsync(ctx->q_p.p, 0xFF, S * 8, ctx->stream));
        CK(cudaMemsetAsync(ctx->q_r2.p, 0xFF, S * 8, ctx->stream));
        CK(cudaMemsetAsync(ctx->q_beta.p, 0xFF, S * 8, ctx->stream));
        CK(cudaMemsetAsync(ctx->q_se.p, 0xFF, S * 8, ctx->stream));
        CK(cudaMemsetAsync(ctx->q_code.p, STOAT_P_NA, S, ctx->stream));
        CK(cudaMemsetAsync(x->q.tail_t.p, 0xFF, S * 8, ctx->stream));  // NaN: no Student-t tail to evaluate
    }
    QuantOut out{ctx->q_allele.as<uint32_t>(), ctx->filtered.as<uint8_t>(), ctx->q_nused.as<uint32_t>(),
                 ctx->q_p.as<double>(),        ctx->q_r2.as<double>(),      ctx->q_beta.as<double>(),
                 ctx->q_se.as<double>(),       ctx->q_code.as<uint8_t>()};
    if (int frc = finalize_timing(ctx)) return frc;
    ctx->timing.launches = 0;
    cudaEvent_t* ev = x->ring[x->run_seq % kRunRing];
    x->ring_kind[x->run_seq % kRunRing] = 2;
    CK(cudaEventRecord(ev[0], ctx->stream));
    if (!S) CK(cudaEventRecord(ev[1], ctx->stream));
    CK(launch_quantitative_impl(matrix_view(ctx), ctx->y.as<double>(), x->q.ys.as<double>(), x->q.sy_all,
                                x->q.syy_all, ctx->n_covar ? ctx->covar.as<double>() : nullptr, uint32_t(ctx->n_covar),
                                ctx->batch, x->max_stage_rows, *thr, out,
                                reinterpret_cast<QuantMoments*>(x->q.mom.p), x->q.part_i.as<uint32_t>(),
                                x->q.part_d.as<double>(), x->q.slow.as<uint32_t>(), x->q.lean_rec.as<uint4>(),
                                x->q.yslices.as<uint8_t>(),
                                x->q.yc.as<double>(), x->q.y_scale, ctx->num_sms, ctx->k1_variant, ctx->stream,
                                &ctx->timing.launches, ev[1], phase, x->q.tail_t.as<double>(),
                                x->q.tail_x.as<double>(), x->q.tail_df.as<uint32_t>(), x->stage_gen, &x->q.lean_key,
                                (ctx->n_covar && quant_covar_fast(uint32_t(ctx->n_covar))) ? x->q.cov_mom.p : nullptr,
                                (ctx->n_covar && quant_covar_fast(uint32_t(ctx->n_covar))) ? x->q.cov_prefix.as<double>() : nullptr,
                                uint32_t(ctx->n_samples_covar)));
    if (phase == 2) CK(cudaEventRecord(ev[1], ctx->stream));
    CK(cudaEventRecord(ev[3], ctx->stream));
    ++x->run_seq;
    x->pending = 2;
    x->q.have_partials = phase == 1;
    ctx->have_quant_result = phase != 1;
    return STOAT_GPU_OK;
}

int stoat_gpu_quantitative_run_staged(stoat_gpu_ctx* ctx, const stoat_gpu_thresholds* thr) { return quant_run(ctx, thr, 0); }

int stoat_gpu_quantitative_moments_staged(stoat_gpu_ctx* ctx, const stoat_gpu_thresholds* thr) {
    const int rc = quant_run(ctx, thr, 1);
    if (rc == STOAT_GPU_ERR_CUDA && ctx && ctx->err.find("not supported") != std::string::npos)
        return fail(ctx, STOAT_GPU_ERR_STATE,
                    "quantitative_moments_staged: the two-path moments kernel does not apply (needs < 65536 local "
                    "haplotypes, no covariates)");
    return rc;
}

int stoat_gpu_quantitative_solve_staged(stoat_gpu_ctx* ctx, const stoat_gpu_thresholds* thr) { return quant_run(ctx, thr, 2); }

int stoat_gpu_quantitative_partials_dev(stoat_gpu_ctx* ctx, uint32_t** ints_dev, uint64_t* n_ints, double** doubles_dev,
                                        uint64_t* n_doubles) {
    if (!ctx || !ints_dev || !n_ints || !doubles_dev || !n_doubles)
        return fail(ctx, STOAT_GPU_ERR_ARG, "quantitative_partials_dev: NULL argument");
    CtxExtra* x = extra_of(ctx);
    if (!x->q.have_partials) return fail(ctx, STOAT_GPU_ERR_STATE, "quantitative_partials_dev: no partial sums (run quantitative_moments_staged)");
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));  // the caller's collective runs on its own stream
    const uint64_t slices = std::max<uint32_t>(quant_lean_slices(ctx->pitch, ctx->n_cols, x->max_stage_rows), 1);
    *ints_dev = x->q.part_i.as<uint32_t>();
    *n_ints = slices * 7 * ctx->batch.n_snarls;
    *doubles_dev = x->q.part_d.as<double>();
    *n_doubles = slices * 3 * ctx->batch.n_snarls;
    return STOAT_GPU_OK;
}


int stoat_gpu_quantitative_fetch(stoat_gpu_ctx* ctx, stoat_gpu_quant_out* out) {
    if (!ctx || !out) return fail(ctx, STOAT_GPU_ERR_ARG, "quantitative_fetch: NULL argument");
    if (!ctx->have_quant_result) return fail(ctx, STOAT_GPU_ERR_STATE, "quantitative_fetch: nothing to fetch");
    CK(cudaSetDevice(ctx->device));
    const uint64_t S = ctx->batch.n_snarls, P = ctx->batch.n_paths;
    const cudaMemcpyKind k = cudaMemcpyDeviceToHost;
    CK(cudaEventRecord(ctx->ev[6], ctx->stream));
    if (P && out->allele_paths) CK(cudaMemcpyAsync(out->allele_paths, ctx->q_allele.p, P * 4, k, ctx->stream));
    if (S && out->filtered) CK(cudaMemcpyAsync(out->filtered, ctx->filtered.p, S, k, ctx->stream));
    if (S && out->n_used) CK(cudaMemcpyAsync(out->n_used, ctx->q_nused.p, S * 4, k, ctx->stream));
    if (S && out->p_value) CK(cudaMemcpyAsync(out->p_value, ctx->q_p.p, S * 8, k, ctx->stream));
    if (S && out->r2) CK(cudaMemcpyAsync(out->r2, ctx->q_r2.p, S * 8, k, ctx->stream));
    if (S && out->beta) CK(cudaMemcpyAsync(out->beta, ctx->q_beta.p, S * 8, k, ctx->stream));
    if (S && out->se) CK(cudaMemcpyAsync(out->se, ctx->q_se.p, S * 8, k, ctx->stream));
    if (S && out->code) CK(cudaMemcpyAsync(out->code, ctx->q_code.p, S, k, ctx->stream));
    CK(cudaEventRecord(ctx->ev[7], ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    if (int rc = finalize_timing(ctx)) return rc;
    CK(cudaEventElapsedTime(&ctx->timing.d2h_ms, ctx->ev[6], ctx->ev[7]));
    return STOAT_GPU_OK;
}

int stoat_gpu_quantitative_batch(stoat_gpu_ctx* ctx, const stoat_gpu_paths* paths, const stoat_gpu_thresholds* thr,
                                 stoat_gpu_quant_out* out) {
    if (!ctx || !paths || !thr || !out) return fail(ctx, STOAT_GPU_ERR_ARG, "quantitative_batch: NULL argument");
    int rc = stoat_gpu_stage_paths(ctx, paths);
    if (rc) return rc;
    const float h2d = ctx->timing.h2d_ms;
    rc = stoat_gpu_quantitative_run_staged(ctx, thr);
    if (rc) return rc;
    rc = stoat_gpu_quantitative_fetch(ctx, out);
    ctx->timing.h2d_ms = h2d;
    if (ctx->batch.n_snarls) ctx->timing.launches += 1;  // prep
    return rc;
}

// ---- binary trait with covariates: logistic regression ---------------------------------------------------
// BinaryCovarSnarlAnalyzer::analyze_and_write_snarl (src/snarl_analyzer.cpp:392-428)
int stoat_gpu_binary_covar_run_staged(stoat_gpu_ctx* ctx, const stoat_gpu_thresholds* thr, int use_covariates) {
    if (!ctx || !thr) return fail(ctx, STOAT_GPU_ERR_ARG, "binary_covar: NULL argument");
    if (!ctx->have_batch) return fail(ctx, STOAT_GPU_ERR_STATE, "binary_covar: no staged batch");
    if (!ctx->have_binary) return fail(ctx, STOAT_GPU_ERR_STATE, "binary_covar: no binary phenotype");
    if (use_covariates && ctx->n_covar && ctx->n_samples_covar * 2 < ctx->n_cols)
        return fail(ctx, STOAT_GPU_ERR_ARG, "binary_covar: fewer covariate rows than samples");
    CK(cudaSetDevice(ctx->device));
    CtxExtra* x = extra_of(ctx);
    LogitState& lg = x->lg;
    const BatchDev& b = ctx->batch;
    const uint64_t S = b.n_snarls;
    CK(lg.counts.reserve(std::max<uint64_t>(S * sizeof(LogitCounts), 16)));
    CK(lg.list.reserve(std::max<uint64_t>(S * 4, 16)));
    CK(lg.n_list.reserve(16));
    CK(lg.allele.reserve(std::max<uint64_t>(b.n_paths * 4, 16)));
    CK(lg.filtered.reserve(std::max<uint64_t>(S, 16)));
    CK(lg.n_used.reserve(std::max<uint64_t>(S * 4, 16)));
    CK(lg.p.reserve(std::max<uint64_t>(S * 8, 16)));
    CK(lg.beta.reserve(std::max<uint64_t>(S * 8, 16)));
    CK(lg.se.reserve(std::max<uint64_t>(S * 8, 16)));
    CK(lg.code.reserve(std::max<uint64_t>(S, 16)));
    if (int frc = finalize_timing(ctx)) return frc;
    ctx->timing.launches = 0;
    LogitOut out{lg.allele.as<uint32_t>(), lg.filtered.as<uint8_t>(), lg.n_used.as<uint32_t>(), lg.p.as<double>(),
                 lg.beta.as<double>(),     lg.se.as<double>(),        lg.code.as<uint8_t>()};
    const bool honour = use_covariates && ctx->n_covar > 0;
    // STOAT_GPU_LOGIT=general sends every snarl through the general kernel (tests compare the two paths)
    const char* mode = std::getenv("STOAT_GPU_LOGIT");
    const bool all_general = honour || ctx->n_cols >= (uint64_t(1) << 32) || (mode && std::string(mode) == "general");
    CK(cudaEventRecord(ctx->ev[2], ctx->stream));
    CK(cudaMemsetAsync(lg.n_list.p, 0, 4, ctx->stream));
    if (!all_general)
        CK(launch_logit_fast(matrix_view(ctx), ctx->case_mask.as<uint8_t>(), b, *thr, lg.counts.as<LogitCounts>(),
                             lg.list.as<uint32_t>(), lg.n_list.as<uint32_t>(), out, ctx->num_sms, ctx->stream,
                             &ctx->timing.launches));
    CK(cudaEventRecord(ctx->ev[3], ctx->stream));
    if (S) {
        CK(launch_logit_general(matrix_view(ctx), ctx->is_case_dev.as<uint8_t>(), honour ? ctx->covar.as<double>() : nullptr,
                                uint32_t(ctx->n_covar), b, *thr, out, all_general ? nullptr : lg.list.as<uint32_t>(),
                                lg.n_list.as<uint32_t>(), S, ctx->num_sms, ctx->stream));
        ctx->timing.launches += 1;
    }
    CK(cudaEventRecord(ctx->ev[4], ctx->stream));
    lg.have_result = true;
    return STOAT_GPU_OK;
}

int stoat_gpu_binary_covar_fetch(stoat_gpu_ctx* ctx, stoat_gpu_binary_covar_out* out) {
    if (!ctx || !out) return fail(ctx, STOAT_GPU_ERR_ARG, "binary_covar_fetch: NULL argument");
    CtxExtra* x = extra_of(ctx);
    LogitState& lg = x->lg;
    if (!lg.have_result) return fail(ctx, STOAT_GPU_ERR_STATE, "binary_covar_fetch: nothing to fetch");
    CK(cudaSetDevice(ctx->device));
    const uint64_t S = ctx->batch.n_snarls, P = ctx->batch.n_paths;
    const cudaMemcpyKind k = cudaMemcpyDeviceToHost;
    CK(cudaEventRecord(ctx->ev[6], ctx->stream));
    if (P && out->allele_paths) CK(cudaMemcpyAsync(out->allele_paths, lg.allele.p, P * 4, k, ctx->stream));
    if (S && out->filtered) CK(cudaMemcpyAsync(out->filtered, lg.filtered.p, S, k, ctx->stream));
    if (S && out->n_used) CK(cudaMemcpyAsync(out->n_used, lg.n_used.p, S * 4, k, ctx->stream));
    if (S && out->p_value) CK(cudaMemcpyAsync(out->p_value, lg.p.p, S * 8, k, ctx->stream));
    if (S && out->beta) CK(cudaMemcpyAsync(out->beta, lg.beta.p, S * 8, k, ctx->stream));
    if (S && out->se) CK(cudaMemcpyAsync(out->se, lg.se.p, S * 8, k, ctx->stream));
    if (S && out->code) CK(cudaMemcpyAsync(out->code, lg.code.p, S, k, ctx->stream));
    CK(cudaMemcpyAsync(&lg.n_general, lg.n_list.p, 4, k, ctx->stream));
    CK(cudaEventRecord(ctx->ev[7], ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaEventElapsedTime(&ctx->timing.table_ms, ctx->ev[2], ctx->ev[3]));  // class counts + IRLS on the counts
    CK(cudaEventElapsedTime(&ctx->timing.stats_ms, ctx->ev[3], ctx->ev[4]));  // general kernel
    CK(cudaEventElapsedTime(&ctx->timing.d2h_ms, ctx->ev[6], ctx->ev[7]));
    ctx->timing.n_general = S ? lg.n_general : 0;
    ctx->timing.n_fisher = 0;
    return STOAT_GPU_OK;
}

int stoat_gpu_binary_covar_batch(stoat_gpu_ctx* ctx, const stoat_gpu_paths* paths, const stoat_gpu_thresholds* thr,
                                 int use_covariates, stoat_gpu_binary_covar_out* out) {
    if (!ctx || !paths || !thr || !out) return fail(ctx, STOAT_GPU_ERR_ARG, "binary_covar_batch: NULL argument");
    int rc = stoat_gpu_stage_paths(ctx, paths);
    if (rc) return rc;
    const float h2d = ctx->timing.h2d_ms;
    rc = stoat_gpu_binary_covar_run_staged(ctx, thr, use_covariates);
    if (rc) return rc;
    rc = stoat_gpu_binary_covar_fetch(ctx, out);
    ctx->timing.h2d_ms = h2d;
    if (ctx->batch.n_snarls) ctx->timing.launches += 1;  // prep
    return rc;
}

// ---- eQTL ------------------------------------------------------------------------------------------
// shared front half: stage the paths, upload the expression matrix and the gene lists
static int eqtl_stage(stoat_gpu_ctx* ctx, const stoat_gpu_paths* paths, const stoat_gpu_eqtl_genes* genes,
                      uint64_t* n_pairs_out) {
    if (!genes->expression || genes->n_genes == 0) return fail(ctx, STOAT_GPU_ERR_ARG, "eqtl: no expression matrix");
    if (genes->n_genes >= 0xFFFFFFFFull) return fail(ctx, STOAT_GPU_ERR_ARG, "eqtl: too many genes");
    if (genes->snarl_gene_off && !genes->gene_idx && paths->n_snarls && genes->snarl_gene_off[paths->n_snarls] != 0)
        return fail(ctx, STOAT_GPU_ERR_ARG, "eqtl: snarl_gene_off without gene_idx");
    if (!genes->snarl_gene_off && genes->gene_idx)
        return fail(ctx, STOAT_GPU_ERR_ARG, "eqtl: gene_idx without snarl_gene_off");
    if (ctx->n_covar && ctx->n_samples_covar * 2 < ctx->n_cols)
        return fail(ctx, STOAT_GPU_ERR_ARG, "eqtl: fewer covariate rows than samples");
    int rc = stoat_gpu_stage_paths(ctx, paths);
    if (rc) return rc;
    const uint64_t S = paths->n_snarls, G = genes->n_genes, n = ctx->n_cols / 2;
    uint64_t n_pairs = S * G;
    if (genes->snarl_gene_off) {
        n_pairs = S ? genes->snarl_gene_off[S] : 0;
        for (uint64_t s = 0; s < S; ++s)
            if (genes->snarl_gene_off[s + 1] < genes->snarl_gene_off[s])
                return fail(ctx, STOAT_GPU_ERR_ARG, "eqtl: snarl_gene_off is not monotone");
        for (uint64_t k = 0; k < n_pairs; ++k)
            if (genes->gene_idx[k] >= G) return fail(ctx, STOAT_GPU_ERR_ARG, "eqtl: gene_idx out of range");
    }
    CK(ctx->e_expr.reserve(std::max<uint64_t>(G * n * 8, 16)));
    CK(cudaMemcpyAsync(ctx->e_expr.p, genes->expression, G * n * 8, cudaMemcpyHostToDevice, ctx->stream));
    if (genes->snarl_gene_off) {
        CK(ctx->e_gene_off.reserve((S + 1) * 8));
        CK(ctx->e_gene_idx.reserve(std::max<uint64_t>(n_pairs * 4, 16)));
        CK(cudaMemcpyAsync(ctx->e_gene_off.p, genes->snarl_gene_off, (S + 1) * 8, cudaMemcpyHostToDevice, ctx->stream));
        if (n_pairs)
            CK(cudaMemcpyAsync(ctx->e_gene_idx.p, genes->gene_idx, n_pairs * 4, cudaMemcpyHostToDevice, ctx->stream));
    }
    const BatchDev& b = ctx->batch;
    CK(ctx->q_allele.reserve(std::max<uint64_t>(b.n_paths * 4, 16)));
    CK(ctx->filtered.reserve(std::max<uint64_t>(S, 16)));
    CK(ctx->q_nused.reserve(std::max<uint64_t>(S * 4, 16)));
    *n_pairs_out = n_pairs;
    return STOAT_GPU_OK;
}

static int eqtl_reserve_pairs(stoat_gpu_ctx* ctx, uint64_t n) {
    CK(ctx->q_p.reserve(std::max<uint64_t>(n * 8, 16)));
    CK(ctx->q_r2.reserve(std::max<uint64_t>(n * 8, 16)));
    CK(ctx->q_beta.reserve(std::max<uint64_t>(n * 8, 16)));
    CK(ctx->q_se.reserve(std::max<uint64_t>(n * 8, 16)));
    CK(ctx->q_code.reserve(std::max<uint64_t>(n, 16)));
    return STOAT_GPU_OK;
}

static QuantOut eqtl_out_view(stoat_gpu_ctx* ctx) {
    return QuantOut{ctx->q_allele.as<uint32_t>(), ctx->filtered.as<uint8_t>(), ctx->q_nused.as<uint32_t>(),
                    ctx->q_p.as<double>(),        ctx->q_r2.as<double>(),      ctx->q_beta.as<double>(),
                    ctx->q_se.as<double>(),       ctx->q_code.as<uint8_t>()};
}

int stoat_gpu_eqtl_batch(stoat_gpu_ctx* ctx, const stoat_gpu_paths* paths, const stoat_gpu_eqtl_genes* genes,
                         const stoat_gpu_thresholds* thr, stoat_gpu_eqtl_out* out) {
    if (!ctx || !paths || !genes || !thr || !out) return fail(ctx, STOAT_GPU_ERR_ARG, "eqtl_batch: NULL argument");
    if (!ctx->have_matrix) return fail(ctx, STOAT_GPU_ERR_STATE, "eqtl_batch: set the matrix first");
    CK(cudaSetDevice(ctx->device));
    uint64_t n_pairs = 0;
    int rc = eqtl_stage(ctx, paths, genes, &n_pairs);
    if (rc) return rc;
    const float h2d = ctx->timing.h2d_ms;
    ctx->have_quant_result = false;  // the pair outputs live in the quantitative result buffers
    if ((rc = eqtl_reserve_pairs(ctx, n_pairs))) return rc;
    const uint64_t S = ctx->batch.n_snarls, P = ctx->batch.n_paths;
    const bool cov_fast = quant_covar_fast(ctx->n_covar ? uint32_t(ctx->n_covar) : 1u);  // (also without covariates: Q = 0)
    if (cov_fast) CK(extra_of(ctx)->q.cov_flag.reserve(std::max<uint64_t>(S, 16)));
    CK(cudaEventRecord(ctx->ev[2], ctx->stream));
    CK(launch_eqtl_general(matrix_view(ctx), ctx->e_expr.as<double>(), uint32_t(genes->n_genes),
                           genes->snarl_gene_off ? ctx->e_gene_off.as<uint64_t>() : nullptr,
                           genes->snarl_gene_off ? ctx->e_gene_idx.as<uint32_t>() : nullptr,
                           ctx->n_covar ? ctx->covar.as<double>() : nullptr, uint32_t(ctx->n_covar), ctx->batch, 0, S, 0,
                           *thr, eqtl_out_view(ctx), ctx->num_sms, ctx->stream, nullptr,
                           (cov_fast && ctx->n_covar) ? extra_of(ctx)->q.cov_prefix.as<double>() : nullptr,
                           uint32_t(ctx->n_samples_covar), cov_fast ? extra_of(ctx)->q.cov_flag.as<uint8_t>() : nullptr, 0));
    CK(cudaEventRecord(ctx->ev[3], ctx->stream));
    const cudaMemcpyKind k = cudaMemcpyDeviceToHost;
    CK(cudaEventRecord(ctx->ev[6], ctx->stream));
    if (P && out->allele_paths) CK(cudaMemcpyAsync(out->allele_paths, ctx->q_allele.p, P * 4, k, ctx->stream));
    if (S && out->filtered) CK(cudaMemcpyAsync(out->filtered, ctx->filtered.p, S, k, ctx->stream));
    if (S && out->n_used) CK(cudaMemcpyAsync(out->n_used, ctx->q_nused.p, S * 4, k, ctx->stream));
    if (n_pairs && out->p_value) CK(cudaMemcpyAsync(out->p_value, ctx->q_p.p, n_pairs * 8, k, ctx->stream));
    if (n_pairs && out->r2) CK(cudaMemcpyAsync(out->r2, ctx->q_r2.p, n_pairs * 8, k, ctx->stream));
    if (n_pairs && out->beta) CK(cudaMemcpyAsync(out->beta, ctx->q_beta.p, n_pairs * 8, k, ctx->stream));
    if (n_pairs && out->se) CK(cudaMemcpyAsync(out->se, ctx->q_se.p, n_pairs * 8, k, ctx->stream));
    if (n_pairs && out->code) CK(cudaMemcpyAsync(out->code, ctx->q_code.p, n_pairs, k, ctx->stream));
    CK(cudaEventRecord(ctx->ev[7], ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaEventElapsedTime(&ctx->timing.table_ms, ctx->ev[2], ctx->ev[3]));
    CK(cudaEventElapsedTime(&ctx->timing.d2h_ms, ctx->ev[6], ctx->ev[7]));
    ctx->timing.h2d_ms = h2d;
    ctx->timing.stats_ms = 0.f;
    ctx->timing.launches = S ? 2 : 0;
    ctx->timing.n_general = uint32_t(S);
    return STOAT_GPU_OK;
}

// ---- dense eQTL scan -------------------------------------------------------------------------------
// |t| below which the two-sided NORMAL tail is already >= p_threshold; Student's t has the heavier tail for every df,
// so such a pair can not have p < p_threshold and its exact tail need not be evaluated.
static double eqtl_z_threshold(double p_threshold) {
    if (!(p_threshold > 0.0)) return 1e300;  // nothing is below
    if (p_threshold > 1.0) return -1.0;      // everything is
    double lo = 0.0, hi = 40.0;              // erfc(40 / sqrt 2) underflows: p below ~1e-350 is never reached
    for (int i = 0; i < 200; ++i) {
        const double mid = 0.5 * (lo + hi);
        if (std::erfc(mid * 0.7071067811865476) > p_threshold) lo = mid; else hi = mid;
    }
    return lo * (1.0 - 1e-9);
}

// centre + pad the expression matrix that already sits in ctx->e_expr
static int eqtl_prepare_expression(stoat_gpu_ctx* ctx, uint64_t G) {
    const uint64_t n = ctx->n_cols / 2, n_pad = (ctx->pitch / 16) * uint64_t(kEqKT);
    CK(ctx->e_yc.reserve(std::max<uint64_t>(G * n_pad * 8, 16)));
    CK(ctx->e_tot.reserve(std::max<uint64_t>(G * 8, 16)));
    CK(ctx->e_tot2.reserve(std::max<uint64_t>(G * 8, 16)));
    CK(launch_eqtl_centre(ctx->e_expr.as<double>(), uint32_t(G), uint32_t(n), uint32_t(n_pad), ctx->e_yc.as<double>(),
                          ctx->e_tot.as<double>(), ctx->e_tot2.as<double>(), ctx->stream));
    if (ctx->eqtl_variant == 0) {
        // the limbs of the centred expression as the shared-memory images of the tcgen05 kernel's B tiles (once per matrix)
        CtxExtra* x = extra_of(ctx);
        const uint64_t g_tiles = (G + kUmTileG - 1) / kUmTileG, n_kb = n_pad / kUmKB;
        CK(x->um_scale.reserve(std::max<uint64_t>(G * 8, 16)));
        CK(x->um_b.reserve(std::max<uint64_t>(g_tiles * n_kb * kUmBBytes, 16)));
        CK(launch_eqtl_umma_prepare_expression(ctx->e_yc.as<double>(), uint32_t(G), uint32_t(n_pad), x->um_scale.as<double>(),
                                               x->um_b.as<uint8_t>(), ctx->stream));
    }
    ctx->e_n_genes = G;
    extra_of(ctx)->expr_cols = ctx->n_cols;
    ctx->have_expression = true;
    ctx->have_eqtl_summary = false;
    return STOAT_GPU_OK;
}

static int eqtl_check_expression_args(stoat_gpu_ctx* ctx, const void* expr, uint64_t n_genes, const char* who) {
    if (!ctx || !expr || n_genes == 0) return fail(ctx, STOAT_GPU_ERR_ARG, std::string(who) + ": no expression matrix");
    if (!ctx->have_matrix) return fail(ctx, STOAT_GPU_ERR_STATE, std::string(who) + ": set the matrix first");
    if (n_genes >= 0xFFFFFFFFull) return fail(ctx, STOAT_GPU_ERR_ARG, std::string(who) + ": too many genes");
    return STOAT_GPU_OK;
}

int stoat_gpu_set_expression(stoat_gpu_ctx* ctx, const double* expression, uint64_t n_genes) {
    if (int rc = eqtl_check_expression_args(ctx, expression, n_genes, "set_expression")) return rc;
    CK(cudaSetDevice(ctx->device));
    const uint64_t n = ctx->n_cols / 2;
    CK(ctx->e_expr.reserve(std::max<uint64_t>(n_genes * n * 8, 16)));
    CK(cudaMemcpyAsync(ctx->e_expr.p, expression, n_genes * n * 8, cudaMemcpyHostToDevice, ctx->stream));
    if (int rc = eqtl_prepare_expression(ctx, n_genes)) return rc;
    CK(cudaStreamSynchronize(ctx->stream));
    return STOAT_GPU_OK;
}

int stoat_gpu_set_expression_dev(stoat_gpu_ctx* ctx, const double* expression_dev, uint64_t n_genes) {
    if (int rc = eqtl_check_expression_args(ctx, expression_dev, n_genes, "set_expression_dev")) return rc;
    CK(cudaSetDevice(ctx->device));
    const uint64_t n = ctx->n_cols / 2;
    CK(ctx->e_expr.reserve(std::max<uint64_t>(n_genes * n * 8, 16)));
    CK(cudaMemcpyAsync(ctx->e_expr.p, expression_dev, n_genes * n * 8, cudaMemcpyDeviceToDevice, ctx->stream));
    if (int rc = eqtl_prepare_expression(ctx, n_genes)) return rc;
    CK(cudaStreamSynchronize(ctx->stream));
    return STOAT_GPU_OK;
}

// the scan over the staged batch and the resident expression matrix; results stay on the device
int stoat_gpu_eqtl_dense_run_staged(stoat_gpu_ctx* ctx, const stoat_gpu_thresholds* thr, double p_threshold) {
    if (!ctx || !thr) return fail(ctx, STOAT_GPU_ERR_ARG, "eqtl_dense_run_staged: NULL argument");
    if (!ctx->have_matrix || !ctx->have_batch || !ctx->have_expression)
        return fail(ctx, STOAT_GPU_ERR_STATE, "eqtl_dense_run_staged: needs the matrix, staged paths and the expression matrix");
    if (ctx->n_covar && ctx->n_samples_covar * 2 < ctx->n_cols)
        return fail(ctx, STOAT_GPU_ERR_ARG, "eqtl: fewer covariate rows than samples");
    CK(cudaSetDevice(ctx->device));
    if (extra_of(ctx)->expr_cols != ctx->n_cols)
        return fail(ctx, STOAT_GPU_ERR_STATE, "eqtl_dense_run_staged: the expression matrix was prepared for another matrix width");
    if (int frc = finalize_timing(ctx)) return frc;
    ctx->have_quant_result = false;  // filtered / allele_paths / pair outputs are shared with the quantitative trait
    ctx->have_eqtl_summary = false;
    const BatchDev& b = ctx->batch;
    const uint64_t S = b.n_snarls, G = ctx->e_n_genes, n = ctx->n_cols / 2;
    CK(ctx->q_allele.reserve(std::max<uint64_t>(b.n_paths * 4, 16)));
    CK(ctx->filtered.reserve(std::max<uint64_t>(S, 16)));
    CK(ctx->q_nused.reserve(std::max<uint64_t>(S * 4, 16)));
    CK(ctx->e_min_p.reserve(std::max<uint64_t>(S * 8, 16)));
    CK(ctx->e_argmin.reserve(std::max<uint64_t>(S * 4, 16)));
    CK(ctx->e_nbelow.reserve(std::max<uint64_t>(S * 4, 16)));
    EqtlDense e{};
    e.yc = ctx->e_yc.as<double>();
    e.tot = ctx->e_tot.as<double>();
    e.tot2 = ctx->e_tot2.as<double>();
    e.n_genes = uint32_t(G);
    e.n_samples = uint32_t(n);
    e.n_kst = uint32_t(ctx->pitch / 16);
    e.n_pad = e.n_kst * kEqKT;
    e.n_snarls = S;
    e.s_pad = (S + kEqTileS - 1) / kEqTileS * kEqTileS;
    // a gene group's rows of yc (~48 MB) stay L2 resident while every snarl tile passes over them
    const uint64_t row_bytes = uint64_t(e.n_pad) * 8;
    // (the tcgen05 form streams 16 KB of A per 32 KB of B and keeps a panel of A images resident too: smaller groups,
    // measured at C4: 256 / 576 / 1152 / 2304 genes per group 397 / 487 / 518 / 533 ms)
    const bool umma_form = ctx->eqtl_variant == 0 && n < 65536;
    uint64_t gpg = std::max<uint64_t>(1, (uint64_t(umma_form ? 10 : 48) << 20) / row_bytes / kEqTileG) * kEqTileG;
    if (const char* v = std::getenv("STOAT_GPU_EQTL_GROUP")) gpg = std::max<uint64_t>(1, uint64_t(std::atoll(v)) / kEqTileG) * kEqTileG;
    e.genes_per_group = uint32_t(std::min<uint64_t>(gpg, (G + kEqTileG - 1) / kEqTileG * kEqTileG));
    e.n_groups = uint32_t((G + e.genes_per_group - 1) / e.genes_per_group);
    CK(ctx->e_dcode.reserve(std::max<uint64_t>(uint64_t(e.n_kst) * e.s_pad * 16, 16)));
    CK(ctx->e_rec.reserve(std::max<uint64_t>(e.s_pad * sizeof(EqtlRec), 16)));
    CK(ctx->e_fb_list.reserve(std::max<uint64_t>(e.s_pad * 4, 16)));
    CK(ctx->e_fb_count.reserve(16));
    CK(ctx->e_slot.reserve(std::max<uint64_t>(e.s_pad * 4, 16)));
    CK(ctx->e_part_key.reserve(std::max<uint64_t>(uint64_t(e.n_groups) * e.s_pad * 8, 16)));
    CK(ctx->e_part_gene.reserve(std::max<uint64_t>(uint64_t(e.n_groups) * e.s_pad * 4, 16)));
    CK(ctx->e_part_cnt.reserve(std::max<uint64_t>(uint64_t(e.n_groups) * e.s_pad * 4, 16)));
    e.dcode = ctx->e_dcode.as<uint8_t>();
    e.rec = ctx->e_rec.as<EqtlRec>();
    e.fb_list = ctx->e_fb_list.as<uint32_t>();
    e.fb_count = ctx->e_fb_count.as<uint32_t>();
    e.dense_count = e.fb_count + 1;
    e.slot_snarl = ctx->e_slot.as<uint32_t>();
    e.part_key = ctx->e_part_key.as<double>();
    e.part_gene = ctx->e_part_gene.as<uint32_t>();
    e.part_cnt = ctx->e_part_cnt.as<uint32_t>();
    e.p_threshold = p_threshold;
    e.z_threshold = eqtl_z_threshold(p_threshold);
    e.min_p = ctx->e_min_p.as<double>();
    e.argmin = ctx->e_argmin.as<uint32_t>();
    e.n_below = ctx->e_nbelow.as<uint32_t>();
    const bool allow_dense = ctx->eqtl_variant != 1 && ctx->n_covar == 0;
    const bool use_umma = ctx->eqtl_variant == 0 && n < 65536;
    EqtlUmma um{};
    if (allow_dense && use_umma) {
        CtxExtra* x = extra_of(ctx);
        um.n_kb = e.n_pad / kUmKB;
        CK(x->um_a.reserve(std::max<uint64_t>((e.s_pad / kUmTileS) * uint64_t(um.n_kb) * kUmABytes, 16)));
        um.a_img = x->um_a.as<uint8_t>();
        um.b_img = x->um_b.as<uint8_t>();
        um.scale = x->um_scale.as<double>();
        um.probe_out = nullptr;
        um.panel = 64;
        um.split = 0;   // measured at C4: one copy per operand image 211 ms, 24 copies of 2 KB + rotated K order 239 ms
        um.rotate = 0;
        if (const char* v = std::getenv("STOAT_GPU_EQTL_SPLIT")) um.split = uint32_t(std::atoi(v));
        if (const char* v = std::getenv("STOAT_GPU_EQTL_ROTATE")) um.rotate = uint32_t(std::atoi(v));
        if (const char* v = std::getenv("STOAT_GPU_EQTL_PANEL")) um.panel = uint32_t(std::max(1, std::atoi(v)));
    }

    CK(cudaEventRecord(ctx->ev[2], ctx->stream));
    CK(cudaMemsetAsync(e.fb_count, 0, 8, ctx->stream));  // general-path list length, dense slot count
    CK(launch_eqtl_dosage_codes(matrix_view(ctx), b, *thr, allow_dense, e, eqtl_out_view(ctx), ctx->stream));
    uint32_t launches = 1;
    CK(cudaEventRecord(ctx->ev[5], ctx->stream));
    if (allow_dense && S) {
        if (use_umma) {
            CK(launch_eqtl_dense_umma(e, um, ctx->stream));
            ++launches;  // the expansion of the dosage codes into the A image
        } else {
            CK(launch_eqtl_dense(e, ctx->num_sms, ctx->stream));
        }
        CK(cudaEventRecord(ctx->ev[6], ctx->stream));
        CK(launch_eqtl_dense_finish(e, ctx->stream));
        launches += 2;
    } else {
        CK(cudaEventRecord(ctx->ev[6], ctx->stream));
    }
    // snarls the dense path does not take (other shapes, covariates): the general kernel, a chunk of pairs at a time
    uint32_t n_fb = 0;
    CK(cudaMemcpyAsync(&n_fb, e.fb_count, 4, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    if (n_fb) {
        const uint64_t chunk = std::max<uint64_t>(1, std::min<uint64_t>(n_fb, (uint64_t(1) << 24) / G + 1));
        if (int rc = eqtl_reserve_pairs(ctx, chunk * G)) return rc;
        const bool cov_fast = quant_covar_fast(ctx->n_covar ? uint32_t(ctx->n_covar) : 1u);
        if (cov_fast) CK(extra_of(ctx)->q.cov_flag.reserve(std::max<uint64_t>(chunk, 16)));
        for (uint64_t i0 = 0; i0 < n_fb; i0 += chunk) {
            const uint64_t ns = std::min<uint64_t>(chunk, n_fb - i0);
            CK(launch_eqtl_general(matrix_view(ctx), ctx->e_expr.as<double>(), uint32_t(G), nullptr, nullptr,
                                   ctx->n_covar ? ctx->covar.as<double>() : nullptr, uint32_t(ctx->n_covar), b, 0, ns, 0,
                                   *thr, eqtl_out_view(ctx), ctx->num_sms, ctx->stream, e.fb_list + i0,
                                   (cov_fast && ctx->n_covar) ? extra_of(ctx)->q.cov_prefix.as<double>() : nullptr,
                                   uint32_t(ctx->n_samples_covar), cov_fast ? extra_of(ctx)->q.cov_flag.as<uint8_t>() : nullptr, G));
            CK(launch_eqtl_summary(ctx->q_p.as<double>(), ctx->q_code.as<uint8_t>(), ctx->filtered.as<uint8_t>(), 0, ns,
                                   uint32_t(G), p_threshold, e.min_p, e.argmin, e.n_below, ctx->stream, e.fb_list + i0));
            launches += 2;
        }
    }
    CK(cudaEventRecord(ctx->ev[3], ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    float all_ms = 0.f;
    CK(cudaEventElapsedTime(&ctx->timing.table_ms, ctx->ev[5], ctx->ev[6]));  // E1, the tensor-core contraction + epilogue
    CK(cudaEventElapsedTime(&all_ms, ctx->ev[2], ctx->ev[3]));
    ctx->timing.stats_ms = all_ms - ctx->timing.table_ms;                     // E0, E2 and the general path for the rest
    ctx->timing.launches = launches;
    ctx->timing.n_general = n_fb;
    ctx->have_eqtl_summary = true;
    return STOAT_GPU_OK;
}

// Test hook of the tcgen05 kernel's mechanics (operand images, descriptors, TMEM round trip): C = A * B^T for int8
// A [128][K], B [n_b_rows][K] (n_b_rows a multiple of 256, K of 128), C int32 [128][n_b_rows], through the SAME kernel the
// scan uses (its epilogue dumps the raw accumulators).  Host buffers.
int stoat_gpu_debug_umma_probe(stoat_gpu_ctx* ctx, const int8_t* a, const int8_t* b, uint32_t n_b_rows, uint32_t k,
                               int32_t* c) {
    if (!ctx || !a || !b || !c) return fail(ctx, STOAT_GPU_ERR_ARG, "umma_probe: NULL argument");
    if (n_b_rows == 0 || n_b_rows % kUmTileN || k == 0 || k % kUmKB) return fail(ctx, STOAT_GPU_ERR_ARG, "umma_probe: bad shape");
    CK(cudaSetDevice(ctx->device));
    CtxExtra* x = extra_of(ctx);
    const uint32_t n_kb = k / kUmKB, n_gt = n_b_rows / kUmTileN;
    std::vector<uint8_t> ai(size_t(n_kb) * kUmABytes), bi(size_t(n_gt) * n_kb * kUmBBytes);
    auto swz = [](uint32_t r, uint32_t kk) { return r * 128u + ((((kk >> 4) ^ (r & 7u)) << 4) | (kk & 15u)); };
    for (uint32_t r = 0; r < uint32_t(kUmTileS); ++r)
        for (uint32_t kk = 0; kk < k; ++kk) ai[size_t(kk / kUmKB) * kUmABytes + swz(r, kk % kUmKB)] = uint8_t(a[size_t(r) * k + kk]);
    for (uint32_t r = 0; r < n_b_rows; ++r)
        for (uint32_t kk = 0; kk < k; ++kk)
            bi[(size_t(r / kUmTileN) * n_kb + kk / kUmKB) * kUmBBytes + swz(r % kUmTileN, kk % kUmKB)] = uint8_t(b[size_t(r) * k + kk]);
    const size_t c_bytes = size_t(kUmTileS) * n_b_rows * 4;
    CK(x->um_probe.reserve(ai.size() + bi.size() + c_bytes + 64));
    uint8_t* d_a = x->um_probe.as<uint8_t>();
    uint8_t* d_b = d_a + ai.size();
    int32_t* d_c = reinterpret_cast<int32_t*>(d_b + bi.size());
    CK(cudaMemcpyAsync(d_a, ai.data(), ai.size(), cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(d_b, bi.data(), bi.size(), cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemsetAsync(d_c, 0xFF, c_bytes, ctx->stream));
    EqtlDense e{};
    e.n_genes = n_b_rows / 8;
    e.genes_per_group = e.n_genes;
    e.n_groups = 1;
    e.s_pad = kUmTileS;
    e.n_snarls = kUmTileS;
    EqtlUmma um{};
    um.a_img = d_a;
    um.b_img = d_b;
    um.n_kb = n_kb;
    um.probe_out = d_c;
    CK(launch_eqtl_dense_umma(e, um, ctx->stream));
    CK(cudaMemcpyAsync(c, d_c, c_bytes, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return STOAT_GPU_OK;
}

int stoat_gpu_eqtl_summary_fetch(stoat_gpu_ctx* ctx, stoat_gpu_eqtl_summary* out) {
    if (!ctx || !out) return fail(ctx, STOAT_GPU_ERR_ARG, "eqtl_summary_fetch: NULL argument");
    if (!ctx->have_eqtl_summary) return fail(ctx, STOAT_GPU_ERR_STATE, "eqtl_summary_fetch: no scan has run");
    CK(cudaSetDevice(ctx->device));
    const uint64_t S = ctx->batch.n_snarls;
    const cudaMemcpyKind k = cudaMemcpyDeviceToHost;
    if (S && out->min_p) CK(cudaMemcpyAsync(out->min_p, ctx->e_min_p.p, S * 8, k, ctx->stream));
    if (S && out->argmin_gene) CK(cudaMemcpyAsync(out->argmin_gene, ctx->e_argmin.p, S * 4, k, ctx->stream));
    if (S && out->n_below) CK(cudaMemcpyAsync(out->n_below, ctx->e_nbelow.p, S * 4, k, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return STOAT_GPU_OK;
}

// the drop-in call with host buffers: stage + upload + scan + fetch
int stoat_gpu_eqtl_dense_summary(stoat_gpu_ctx* ctx, const stoat_gpu_paths* paths, const stoat_gpu_eqtl_genes* genes,
                                 const stoat_gpu_thresholds* thr, double p_threshold, stoat_gpu_eqtl_summary* out) {
    if (!ctx || !paths || !genes || !thr || !out) return fail(ctx, STOAT_GPU_ERR_ARG, "eqtl_dense_summary: NULL argument");
    if (!ctx->have_matrix) return fail(ctx, STOAT_GPU_ERR_STATE, "eqtl_dense_summary: set the matrix first");
    if (genes->snarl_gene_off || genes->gene_idx)
        return fail(ctx, STOAT_GPU_ERR_ARG, "eqtl_dense_summary: dense scan only (snarl_gene_off / gene_idx must be NULL)");
    CK(cudaSetDevice(ctx->device));
    uint64_t n_pairs = 0;
    int rc = eqtl_stage(ctx, paths, genes, &n_pairs);
    if (rc) return rc;
    if ((rc = eqtl_prepare_expression(ctx, genes->n_genes))) return rc;
    if ((rc = stoat_gpu_eqtl_dense_run_staged(ctx, thr, p_threshold))) return rc;
    return stoat_gpu_eqtl_summary_fetch(ctx, out);
}

// Text of a value column — stoat::set_precision (src/utils.cpp:5-15) and the literal strings of `codes`, formatted on the
// device (kernels_post.cu); the few values whose decimal rounding FP64 cannot decide are redone here with the C library,
// so slot i always holds exactly what stoat_gpu_format_value(values[i], codes[i]) would write.
int stoat_gpu_format_values(stoat_gpu_ctx* ctx, const double* values, const uint8_t* codes, uint64_t n, char* slots,
                            uint8_t* lengths) {
    if (!ctx) return STOAT_GPU_ERR_ARG;
    if (n == 0) return STOAT_GPU_OK;
    if (!values || !slots || !lengths) return fail(ctx, STOAT_GPU_ERR_ARG, "format_values: NULL argument");
    CK(cudaSetDevice(ctx->device));
    CtxExtra* x = extra_of(ctx);
    // scratch: slots [n][16] (16-byte stores) | values [n] f64 | codes [n] | lengths [n] | unsure [n]
    const size_t off_slots = 0, off_values = size_t(n) * 16, off_codes = off_values + size_t(n) * 8, off_len = off_codes + n,
                 off_unsure = off_len + n;
    CK(x->fmt.reserve(off_unsure + n + 64));
    uint8_t* base = x->fmt.as<uint8_t>();
    CK(cudaMemcpyAsync(base + off_values, values, n * 8, cudaMemcpyHostToDevice, ctx->stream));
    if (codes) CK(cudaMemcpyAsync(base + off_codes, codes, n, cudaMemcpyHostToDevice, ctx->stream));
    CK(launch_format_values(reinterpret_cast<const double*>(base + off_values), codes ? base + off_codes : nullptr, n,
                            reinterpret_cast<char*>(base + off_slots), base + off_len, base + off_unsure, ctx->stream));
    x->fmt_unsure.resize(n);
    CK(cudaMemcpyAsync(slots, base + off_slots, n * 16, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaMemcpyAsync(lengths, base + off_len, n, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaMemcpyAsync(x->fmt_unsure.data(), base + off_unsure, n, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    for (uint64_t i = 0; i < n; ++i)
        if (x->fmt_unsure[i]) {
            char buf[32];
            const int l = stoat_gpu_format_value(values[i], codes ? codes[i] : uint8_t(STOAT_P_VALUE), buf, sizeof buf);
            if (l < 0 || l > 15) return fail(ctx, STOAT_GPU_ERR_ARG, "format_values: value does not fit a slot");
            std::memset(slots + i * 16, 0, 16);
            std::memcpy(slots + i * 16, buf, size_t(l));
            lengths[i] = uint8_t(l);
        }
    return STOAT_GPU_OK;
}

// set_precision (src/utils.cpp:5-15): iostream scientific / defaultfloat at precision 4 are "%.4e" / "%.4g"
int stoat_gpu_format_value(double value, uint8_t code, char* buf, size_t buf_len) {
    if (!buf || buf_len == 0) return -1;
    int n;
    switch (code) {
        case STOAT_P_NA: n = std::snprintf(buf, buf_len, "NA"); break;
        case STOAT_P_ZERO: n = std::snprintf(buf, buf_len, "0"); break;
        case STOAT_P_ONE: n = std::snprintf(buf, buf_len, "1"); break;
        case STOAT_P_DBLMAX: n = std::snprintf(buf, buf_len, "%.4g", 1.7976931348623157e308); break;
        default:
            if (std::fabs(value) < 1e-1 && value != 0.0)
                n = std::snprintf(buf, buf_len, "%.4e", value);
            else
                n = std::snprintf(buf, buf_len, "%.4g", value);
    }
    return (n < 0 || size_t(n) >= buf_len) ? -1 : n;
}

int stoat_gpu_last_timing(const stoat_gpu_ctx* ctx, stoat_gpu_timing* out) {
    if (!ctx || !out) return STOAT_GPU_ERR_ARG;
    if (int rc = finalize_timing(const_cast<stoat_gpu_ctx*>(ctx))) return rc;  // waits for a run that is still in flight
    *out = ctx->timing;
    return STOAT_GPU_OK;
}

int stoat_gpu_mean_timing(stoat_gpu_ctx* ctx, uint32_t n_runs, stoat_gpu_timing* out) {
    if (!ctx || !out) return STOAT_GPU_ERR_ARG;
    if (int rc = finalize_timing(ctx)) return rc;
    CtxExtra* x = extra_of(ctx);
    *out = ctx->timing;
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    uint64_t n = std::min<uint64_t>(std::min<uint64_t>(n_runs, x->run_seq), kRunRing);
    if (n == 0) return STOAT_GPU_OK;
    const uint8_t kind = x->ring_kind[(x->run_seq - 1) % kRunRing];
    double table = 0.0, stats = 0.0;
    uint64_t used = 0;
    for (uint64_t i = 0; i < n; ++i) {
        const uint64_t slot = (x->run_seq - 1 - i) % kRunRing;
        if (x->ring_kind[slot] != kind) break;  // an older run of the other trait
        cudaEvent_t* ev = x->ring[slot];
        float t = 0.f, u = 0.f;
        CK(cudaEventElapsedTime(&t, ev[0], ev[1]));
        CK(cudaEventElapsedTime(&u, kind == 1 ? ev[2] : ev[1], ev[3]));
        table += t;
        stats += u;
        ++used;
    }
    out->table_ms = float(table / double(used));
    out->stats_ms = float(stats / double(used));
    return STOAT_GPU_OK;
}

}  // extern "C"
