// ctx.hpp — the context behind the C ABI (include/stoat_gpu.h) and the launch interfaces of the
// kernel translation units.
#pragma once
#include <cstdint>
#include <cstdio>
#include <string>
#include <cuda_runtime.h>

#include "../../include/stoat_gpu.h"

namespace stoat {

// ---- device buffer that only grows --------------------------------------------------------------
struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    cudaError_t reserve(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        // The old contents are never needed after a growth, so the old block goes first (a C5-sized matrix and its
        // replacement do not fit side by side); on failure the buffer is left empty, never dangling.
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
        // grow geometrically a little so repeated batches of similar size do not realloc
        size_t want = bytes + bytes / 8 + 256;
        cudaError_t e = cudaMalloc(&p, want);
        if (e != cudaSuccess) {
            (void)cudaGetLastError();  // clear the failed attempt: it would surface at the next launch's error check
            p = nullptr;
            e = cudaMalloc(&p, bytes);
            want = bytes;
        }
        if (e == cudaSuccess)
            cap = want;
        else {
            (void)cudaGetLastError();
            p = nullptr;
        }
        return e;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
    template <typename T>
    T* as() const {
        return static_cast<T*>(p);
    }
};

// Per-snarl record produced by the prep kernel and consumed by the TMA producer warp (32 bytes,
// one coalesced 2 x 16-byte load).  rows[] are the snarl's distinct matrix rows.
constexpr int kMaxStageRows = 6;  // distinct rows a snarl may have on the TMA path (else: LDG path)
struct __align__(32) SnarlRec {
    uint32_t rows[kMaxStageRows];  // distinct rows (global row index)
    uint32_t blob16;               // offset of the path program in the blob array, 16-byte units
    uint8_t n_rows;                // number of distinct rows (0 when the snarl is on the LDG path)
    uint8_t blob_len16;            // path-program length in 16-byte units (>= 1)
    uint16_t n_paths;
};
static_assert(sizeof(SnarlRec) == 32, "SnarlRec must be 32 bytes");

constexpr int kMaxBlob16 = 4;  // path programs longer than 64 bytes go to the LDG path

// device view of a staged batch
struct BatchDev {
    uint64_t n_snarls = 0, n_paths = 0, n_edges = 0;
    const uint64_t* snarl_path_off = nullptr;
    const uint64_t* path_edge_off = nullptr;
    const uint32_t* edge_row = nullptr;
    SnarlRec* rec = nullptr;
    uint8_t* blob = nullptr;      // path programs, see kernels_binary.cu
    uint32_t* big_list = nullptr;  // snarls that take the LDG path
    uint32_t* counters = nullptr;  // [0] = number of big snarls, [1] = error flag, [2] = widest stage (rows)
};

struct MatrixDev {
    const uint8_t* rows = nullptr;  // repacked, row pitch bytes each, padding bits zero
    uint64_t n_rows = 0, n_cols = 0, pitch = 0;
};

// ---- kernel launchers (defined in the kernels_*.cu files) ---------------------------------------
// K0
cudaError_t launch_repack_rows(const uint8_t* raw, uint64_t raw_bytes, uint64_t row_begin, uint64_t row_end,
                               uint64_t n_cols, uint64_t pitch, uint8_t* out, cudaStream_t st);
cudaError_t launch_scatter_bits(const uint32_t* row, const uint32_t* col, uint64_t n_pairs, uint64_t n_rows, uint64_t n_cols,
                                uint64_t pitch, uint8_t* out, uint32_t* bad, int num_sms, cudaStream_t st);
cudaError_t launch_gather_rows(const uint8_t* rows, uint64_t pitch, const uint32_t* idx, uint64_t n, uint8_t* out,
                               cudaStream_t st);
cudaError_t launch_build_case_mask(const uint8_t* is_case, uint64_t n_samples, uint64_t pitch, uint8_t* mask,
                                   cudaStream_t st);
// binary
cudaError_t launch_binary_prep(const BatchDev& b, uint64_t n_matrix_rows, cudaStream_t st);
cudaError_t launch_path_and_popc(const MatrixDev& m, const uint8_t* case_mask, uint32_t n_case_hap, const BatchDev& b,
                                 uint32_t n_big, uint32_t max_stage_rows, uint32_t* counts, int num_sms, int variant,
                                 cudaStream_t st, uint32_t* launches, uint64_t s_begin = 0,
                                 uint64_t s_count = ~uint64_t(0), bool lean_registers = false);
cudaError_t launch_binary_stats(const BatchDev& b, const uint32_t* counts, const stoat_gpu_thresholds& thr,
                                uint64_t* total_sum, uint8_t* filtered, double* p_chi2, double* p_fisher,
                                uint8_t* chi2_code, uint8_t* fisher_code, uint4* fisher_tab, uint32_t* fisher_idx,
                                uint32_t* fisher_counters, int num_sms, cudaStream_t st, uint64_t s_begin = 0,
                                uint64_t s_count = ~uint64_t(0), bool beside_count_pass = false);
// post-processing (kernels_post.cu): Benjamini-Hochberg adjustment
size_t bh_temp_bytes(uint64_t n);
cudaError_t launch_bh_adjust(const double* p, uint64_t n, double* out, uint64_t* keys_sorted, uint32_t* idx_in,
                             uint32_t* idx_sorted, double* a, double* b, void* temp, size_t temp_bytes, cudaStream_t st);
// the same with its scratch in a file-scope device array (no allocation; stoat_gpu_bh_adjust_static)
uint64_t bh_static_max_n();
cudaError_t launch_bh_adjust_static(int device, const double* p, uint64_t n, double* out, uint32_t* bad, cudaStream_t st);
// text of the value columns (kernels_post.cu): slots [n][16] NUL padded, len [n], unsure [n] (1: format on the host)
cudaError_t launch_format_values(const double* v, const uint8_t* code, uint64_t n, char* slots, uint8_t* len,
                                 uint8_t* unsure, cudaStream_t st);
// quantitative
struct QuantMoments;  // kernels_quant.cu
struct QuantOut {
    uint32_t* allele_paths;
    uint8_t* filtered;
    uint32_t* n_used;
    double* p;
    double* r2;
    double* beta;
    double* se;
    uint8_t* code;
};
cudaError_t launch_quantitative_impl(const MatrixDev& m, const double* y_raw, const double* ys_perm,
                                     double sy_all, double syy_all, const double* covar, uint32_t n_covar,
                                     const BatchDev& b, uint32_t max_stage_rows, const stoat_gpu_thresholds& thr,
                                     QuantOut out, QuantMoments* mom, uint32_t* part_i, double* part_d,
                                     uint32_t* slow_list, uint4* lean_rec, const uint8_t* yslices, const double* yc,
                                     double y_scale, int num_sms, int variant, cudaStream_t st, uint32_t* launches,
                                     cudaEvent_t ev_table_done,  // recorded after the moments pass (K3M / K3R / K3)
                                     int phase,  // 0: whole step; 1: moments only; 2: solve only (sample-sharded)
                                     double* tail_t, double* tail_x, uint32_t* tail_df,  // [n_snarls] K4 -> K4b hand-over
                                     uint64_t stage_gen = 0,         // identifies the staged paths (rec / blob)
                                     uint64_t* lean_key = nullptr,   // in/out: what lean_rec holds (nullptr: rebuild every step)
                                     void* cov_mom = nullptr,        // covariate runs: [n_snarls] moments (quant_covar_moments_bytes)
                                     const double* cov_prefix = nullptr,  // ... the prefix sums of the covariate rows
                                                                          // [cov_rows + 1][Q + Q (Q + 1) / 2], then the
                                                                          // covariates by column [Q][cov_rows]
                                     uint32_t cov_rows = 0);
size_t quant_covar_moments_bytes(uint64_t n_snarls);
bool quant_covar_fast(uint32_t n_covar);  // the covariate run takes K3C / K4C (quant_covar.cuh) instead of K3G for every snarl
bool quant_fast_path(uint64_t pitch, uint64_t n_hap, uint32_t max_stage_rows, int variant, uint32_t n_covar);  // K4 initialises the outputs
size_t quant_yslices_bytes(uint64_t pitch);  // K3M's phenotype slice table
// slices of the haplotype range the two-path kernels (K3M / K3R) use: sizes the partial-sum buffers (0: neither runs)
uint32_t quant_lean_slices(uint64_t pitch, uint64_t n_hap, uint32_t stage_rows);
// logistic regression (kernels_logit.cu: class-count fast path; kernels_quant.cu: the general kernel's logistic mode)
struct LogitCounts {  // per two-path snarl: haplotypes per path, samples per dosage class, and the cases among them
    uint32_t al0, al1, used, full0, half, third, tthird, c_used, c_full0, c_half, c_third, c_tthird;
};
struct LogitOut {
    uint32_t* allele_paths;
    uint8_t* filtered;
    uint32_t* n_used;
    double* p;
    double* beta;
    double* se;
    uint8_t* code;
};
cudaError_t launch_logit_fast(const MatrixDev& m, const uint8_t* case_mask, const BatchDev& b,
                              const stoat_gpu_thresholds& thr, LogitCounts* counts, uint32_t* general_list,
                              uint32_t* n_general, LogitOut out, int num_sms, cudaStream_t st, uint32_t* launches);
// snarls list[0 .. *n_list_dev) (list == nullptr: every snarl of the batch); n_max bounds the grid
cudaError_t launch_logit_general(const MatrixDev& m, const uint8_t* is_case, const double* covar, uint32_t n_covar,
                                 const BatchDev& b, const stoat_gpu_thresholds& thr, LogitOut out, const uint32_t* list,
                                 const uint32_t* n_list_dev, uint64_t n_max, int num_sms, cudaStream_t st);
cudaError_t launch_centre_permute(const double* y, uint32_t n, double mean, uint32_t n_words, double* yc, double* ys,
                                  cudaStream_t st);
size_t quant_moments_bytes(uint64_t n_snarls);
// eQTL (general path over snarls [s_begin, s_begin + n_snarls); per-pair outputs at pair - item_base)
cudaError_t launch_eqtl_general(const MatrixDev& m, const double* expression, uint32_t n_genes,
                                const uint64_t* snarl_gene_off, const uint32_t* gene_idx, const double* covar,
                                uint32_t n_covar, const BatchDev& b, uint64_t s_begin, uint64_t n_snarls,
                                uint64_t item_base, const stoat_gpu_thresholds& thr, QuantOut out, int num_sms,
                                cudaStream_t st, const uint32_t* list = nullptr,
                                // covariate runs (quant_covar.cuh): prefix tables + column-major covariates, their row
                                // count, a scratch byte per snarl of the call; nullptr: every snarl on the general kernel
                                const double* cov_prefix = nullptr, uint32_t cov_rows = 0, uint8_t* cov_flag = nullptr,
                                uint64_t max_items_per_snarl = 0);
cudaError_t launch_eqtl_summary(const double* p, const uint8_t* code, const uint8_t* filtered, uint64_t s_begin,
                                uint64_t n_snarls, uint32_t n_genes, double p_threshold, double* min_p,
                                uint32_t* argmin, uint32_t* n_below, cudaStream_t st, const uint32_t* list = nullptr);

}  // namespace stoat

// ---- the context ----------------------------------------------------------------------------------
struct stoat_gpu_ctx {
    int device = 0;
    int num_sms = 148;
    cudaStream_t stream = nullptr;
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t ev[8] = {};
    cudaEvent_t chunk_ev[2] = {};
    std::string err;
    stoat_gpu_timing timing = {};
    int k1_variant = 0;  // 0 = TMA pipeline (default), 1 = LDG only (STOAT_GPU_K1=ldg)

    // matrix
    stoat::DevBuf rows;        // repacked
    stoat::DevBuf raw_stage[2];  // upload staging (chunked H2D of the reference-layout bytes)
    uint64_t n_rows = 0, n_cols = 0, pitch = 0;
    bool have_matrix = false;

    // phenotypes
    stoat::DevBuf case_mask, is_case_dev, y, covar;
    uint64_t n_samples_binary = 0, n_samples_quant = 0, n_samples_covar = 0, n_covar = 0;
    uint32_t n_case_hap = 0;
    bool have_binary = false, have_quant = false;

    // staged batch
    stoat::DevBuf d_snarl_path_off, d_path_edge_off, d_edge_row, d_rec, d_blob, d_big, d_counters;
    stoat::BatchDev batch;
    uint32_t n_big = 0;
    bool have_batch = false;

    // results (device)
    stoat::DevBuf counts, total_sum, filtered_bin, p_a, p_b, code_a, code_b;  // binary (its own filter flags)
    stoat::DevBuf filtered;                                                    // quantitative / eQTL filter flags
    stoat::DevBuf fisher_tab, fisher_idx, fisher_cnt;                      // Fisher work list
    stoat::DevBuf q_allele, q_nused, q_p, q_r2, q_beta, q_se, q_code;      // quantitative / eQTL pairs
    stoat::DevBuf e_expr, e_gene_off, e_gene_idx, e_min_p, e_argmin, e_nbelow;  // eQTL
    // dense eQTL scan (kernels_eqtl.cu)
    stoat::DevBuf e_yc, e_tot, e_tot2, e_dcode, e_rec, e_fb_list, e_fb_count, e_slot, e_part_key, e_part_gene, e_part_cnt;
    uint64_t e_n_genes = 0;
    bool have_expression = false, have_eqtl_summary = false;
    int eqtl_variant = 0;  // 0 = dense DMMA path where it applies (default), 1 = general kernel only (STOAT_GPU_EQTL=general)
    bool have_binary_counts = false, have_binary_result = false, have_quant_result = false;
};
