/*
 * stoat_gpu.h — C ABI of the B200 (sm_100a) implementation of STOAT's per-snarl
 * association-testing hot path (`stoat vcf`).
 *
 * The reference (Pa-Tou/stoat) has no plugin / FFI interface: its hot loop is the body of
 *   SnarlAnalyzer::process_snarls_by_chromosome_chunk   (src/snarl_analyzer.cpp:161-172)
 * which calls the virtual analyze_and_write_snarl (src/snarl_analyzer.hpp:66) once per snarl.
 * This ABI replaces that loop body for a whole chromosome chunk with ONE batched call per
 * analyzer type.  Everything is plain pointers and sizes; the caller owns every buffer;
 * the context owns device memory only.  There is no CPU fallback: every entry point fails
 * with STOAT_GPU_ERR_CUDA when no sm_100 device is usable.
 *
 * Data contract (all host pointers unless a name ends in _dev):
 *
 *  presence matrix   the reference's EdgeBySampleMatrix bytes (src/matrix.hpp:51-55,
 *                    src/matrix.cpp:67-84): bit index = row * n_cols + col, LSB-first inside
 *                    each byte, rows NOT byte aligned.  rows = graph edges, cols = haplotypes
 *                    (2 per sample, haplotype h belongs to sample h/2).
 *  snarl paths       CSR produced by the host shim from Snarl_data_t::snarl_paths with the
 *                    reference's decompose_path_to_edges + identify_path front half
 *                    (src/snarl_analyzer.cpp:263-274, 321-337):
 *                      snarl_path_off[S+1]  first path of each snarl
 *                      path_edge_off[P+1]   first checked edge of each path
 *                      edge_row[sum R]      matrix row of each checked edge; edges touching
 *                                           node id 0 are already dropped (:328-330);
 *                                           STOAT_GPU_MISSING_EDGE marks an edge absent from
 *                                           the matrix => the path matches nobody (:334-336).
 *                    A path with zero checked edges matches every haplotype (:343-354).
 *  results           one entry per snarl (or per path), in input order.  P-values come back
 *                    as (double, code) pairs; the code selects the reference's literal strings
 *                    ("NA", "0", "1", set_precision(DBL_MAX)), otherwise the double is
 *                    formatted with set_precision (src/utils.cpp:5-15) by the host shim.
 */
#ifndef STOAT_GPU_H
#define STOAT_GPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define STOAT_GPU_ABI_VERSION 1
#if defined(__GNUC__)
#define STOAT_GPU_API __attribute__((visibility("default")))
#else
#define STOAT_GPU_API
#endif
#define STOAT_GPU_MISSING_EDGE 0xFFFFFFFFu
/* hard limits of the device kernels (a snarl above them is reported, never silently mangled) */
#define STOAT_GPU_MAX_PATHS_PER_SNARL 64   /* paths (table columns) in one snarl            */
#define STOAT_GPU_MAX_OLS_FEATURES 16      /* 1 + dosage columns + covariates              */

/* return codes */
enum {
    STOAT_GPU_OK = 0,
    STOAT_GPU_ERR_ARG = -1,     /* null pointer / inconsistent sizes / limits exceeded      */
    STOAT_GPU_ERR_CUDA = -2,    /* CUDA runtime failure or no sm_100 device                 */
    STOAT_GPU_ERR_STATE = -3,   /* call order (e.g. batch before set_matrix)                */
    STOAT_GPU_ERR_OOM = -4      /* device or pinned-host allocation failed                  */
};

/* how a p-value (or statistic) must be printed; mirrors the string returns of
 * FisherKhi2 (src/stats_test.cpp:189,197,230-234,269,313,323)                              */
enum {
    STOAT_P_VALUE = 0,   /* print set_precision(value)                                      */
    STOAT_P_NA = 1,      /* "NA"                                                            */
    STOAT_P_ZERO = 2,    /* "0"   (Fisher recurrence overflowed, stats_test.cpp:312-314)    */
    STOAT_P_ONE = 3,     /* "1"   (Fisher centre mass is zero, stats_test.cpp:322-324)      */
    STOAT_P_DBLMAX = 4,  /* set_precision(DBL_MAX) (an expected count is zero, :196-197)    */
    STOAT_P_UNDEF = 5,   /* the reference reads out of bounds here (zero dosage columns
                            reach linear_regression, stats_test.cpp:487-489): value is NaN  */
    STOAT_P_OVERLIMIT = 6, /* snarl exceeds STOAT_GPU_MAX_*: not computed (value NaN); the host
                            must run its own per-snarl analyzer for it                      */
    STOAT_P_DEFERRED = 7  /* sample-sharded quantitative run only: the snarl needs every column
                            (not a two-path table, rank deficient, near-perfect fit): replay it
                            with the whole rows (stoat_b200/shard.py does)                   */
};

/* filtering thresholds, defaults of `stoat vcf` (src/subcommand/vcf.cpp:61-73)             */
typedef struct stoat_gpu_thresholds {
    uint64_t min_individuals; /* default 3   */
    uint64_t min_haplotypes;  /* default 5   */
    double maf_threshold;     /* default 0.05 */
} stoat_gpu_thresholds;

/* flattened snarl paths of one batch (see "snarl paths" above)                             */
typedef struct stoat_gpu_paths {
    uint64_t n_snarls;
    const uint64_t* snarl_path_off; /* [n_snarls + 1]                 */
    const uint64_t* path_edge_off;  /* [n_paths + 1], n_paths = snarl_path_off[n_snarls]   */
    const uint32_t* edge_row;       /* [path_edge_off[n_paths]]       */
} stoat_gpu_paths;

/* outputs of the binary analyzer: BinarySnarlAnalyzer::analyze_and_write_snarl
 * (src/snarl_analyzer.cpp:358-390)                                                          */
typedef struct stoat_gpu_binary_out {
    uint32_t* g0;          /* [n_paths] controls on path, BEFORE empty-column removal
                              (create_binary_table, src/binary_table.cpp:19-44)             */
    uint32_t* g1;          /* [n_paths] cases on path                                       */
    uint64_t* total_sum;   /* [n_snarls] sum over paths of matching haplotypes              */
    uint8_t* filtered;     /* [n_snarls] 1 = filtration_binary_table said "filter"
                              (src/snarl_analyzer.cpp:705-732); p-values are then unset     */
    double* p_chi2;        /* [n_snarls] chi2_2x2 / chi2_2xN (src/stats_test.cpp:181-257)   */
    double* p_fisher;      /* [n_snarls] fastFishersExactTest (:264-360), 2 columns only    */
    uint8_t* chi2_code;    /* [n_snarls] STOAT_P_*                                           */
    uint8_t* fisher_code;  /* [n_snarls] STOAT_P_*                                           */
} stoat_gpu_binary_out;

/* outputs of the quantitative analyzer: QuantitativeSnarlAnalyzer::analyze_and_write_snarl
 * (src/snarl_analyzer.cpp:431-467) -> LinearRegression::linear_regression
 * (src/stats_test.cpp:424-506)                                                              */
typedef struct stoat_gpu_quant_out {
    uint32_t* allele_paths; /* [n_paths] haplotypes on each path (0 for empty paths)
                               (process_table_quantitative, src/quantitative_table.cpp:57-58) */
    uint8_t* filtered;      /* [n_snarls] filtration_quantitative_table (:542-579)          */
    uint32_t* n_used;       /* [n_snarls] rows of the regression table (samples with a hit) */
    double* p_value;        /* [n_snarls] two-sided Student-t p of the FIRST dosage column  */
    double* r2;             /* [n_snarls]                                                    */
    double* beta;           /* [n_snarls]                                                    */
    double* se;             /* [n_snarls]                                                    */
    uint8_t* code;          /* [n_snarls] STOAT_P_VALUE or STOAT_P_UNDEF                     */
} stoat_gpu_quant_out;

/* outputs of the binary analyzer with covariates: BinaryCovarSnarlAnalyzer::analyze_and_write_snarl
 * (src/snarl_analyzer.cpp:392-428) -> LogisticRegression::logistic_regression
 * (src/stats_test.cpp:49-176): the dosage table of the quantitative analyzer against the binary
 * phenotype, IRLS with an L2 penalty of 1e-4, Wald test, Holm over the dosage columns.       */
typedef struct stoat_gpu_binary_covar_out {
    uint32_t* allele_paths; /* [n_paths]                                                     */
    uint8_t* filtered;      /* [n_snarls] filtration_quantitative_table (:542-579)          */
    uint32_t* n_used;       /* [n_snarls]                                                    */
    double* p_value;        /* [n_snarls] smallest Holm-adjusted Wald p over the dosage columns */
    double* beta;           /* [n_snarls] coefficient of that column                         */
    double* se;             /* [n_snarls]                                                    */
    uint8_t* code;          /* [n_snarls] STOAT_P_VALUE, STOAT_P_NA ("NA" triple: LDLT failure or
                               no convergence in 100 iterations, :107,120) or STOAT_P_UNDEF  */
} stoat_gpu_binary_covar_out;

/* eQTL: EQTLSnarlAnalyzer::analyze_and_write_snarl (src/snarl_analyzer.cpp:493-539): one OLS
 * per (snarl, gene in window).  Gene windows are resolved on the host (found_gene_snarl,
 * :471-491) into a CSR of gene indices per snarl; pair k of the output is
 * (snarl s, gene gene_idx[k]) for snarl_gene_off[s] <= k < snarl_gene_off[s+1].             */
typedef struct stoat_gpu_eqtl_genes {
    uint64_t n_genes;
    const double* expression;        /* [n_genes * n_samples] row-major, gene major         */
    const uint64_t* snarl_gene_off;  /* [n_snarls + 1]; NULL = dense (every gene, every snarl) */
    const uint32_t* gene_idx;        /* [snarl_gene_off[n_snarls]]; NULL when dense          */
} stoat_gpu_eqtl_genes;

typedef struct stoat_gpu_eqtl_out {
    uint32_t* allele_paths; /* [n_paths]                                                     */
    uint8_t* filtered;      /* [n_snarls]                                                    */
    uint32_t* n_used;       /* [n_snarls]                                                    */
    double* p_value;        /* [n_pairs]                                                     */
    double* r2;             /* [n_pairs]                                                     */
    double* beta;           /* [n_pairs]                                                     */
    double* se;             /* [n_pairs]                                                     */
    uint8_t* code;          /* [n_pairs]                                                     */
} stoat_gpu_eqtl_out;

/* device-side reduction of a dense eQTL scan whose pair table cannot be returned whole
 * (200k snarls x 20k genes = 4e9 rows); mirrors the reference's --table-threshold idea
 * (src/snarl_analyzer.cpp:524-527): per snarl the best gene and the number of significant ones */
typedef struct stoat_gpu_eqtl_summary {
    double* min_p;          /* [n_snarls] smallest p over genes (NaN if filtered)            */
    uint32_t* argmin_gene;  /* [n_snarls]                                                    */
    uint32_t* n_below;      /* [n_snarls] genes with p < p_threshold                         */
} stoat_gpu_eqtl_summary;

typedef struct stoat_gpu_ctx stoat_gpu_ctx;

/* timing of the last batch call, measured with CUDA events on the context's stream          */
typedef struct stoat_gpu_timing {
    float h2d_ms;      /* host -> device copies of the batch inputs                          */
    float repack_ms;   /* K0 repack_rows (only when the matrix was (re)uploaded)              */
    float table_ms;    /* K1 path_and_popc / K3 dosage pass / dense eQTL: the DMMA contraction (E1) */
    float stats_ms;    /* K2 binary_stats / K4 ols_solve / dense eQTL: E0 + E2 + general-path rest  */
    float d2h_ms;      /* device -> host copies of the results                               */
    uint32_t launches; /* kernels launched by the call                                       */
    uint32_t n_general; /* snarls that left the TMA-staged fast path (LDG / general kernels)  */
    uint32_t n_fisher;  /* 2x2 tables that went through the Fisher work list                  */
} stoat_gpu_timing;

/* ---- lifetime -------------------------------------------------------------------------- */
STOAT_GPU_API int stoat_gpu_abi_version(void);
/* one context per host thread and per device (the reference calls the analyzer from OpenMP
 * threads, src/snarl_analyzer.cpp:165; a context is NOT thread safe)                        */
STOAT_GPU_API int stoat_gpu_create(stoat_gpu_ctx** out, int device);
STOAT_GPU_API void stoat_gpu_destroy(stoat_gpu_ctx* ctx);
/* message of the last failure on this context ("" if none); ctx may be NULL for create()    */
STOAT_GPU_API const char* stoat_gpu_last_error(const stoat_gpu_ctx* ctx);

/* ---- per-chromosome state (replaces the const-refs held by SnarlAnalyzer,
 *      src/snarl_analyzer.hpp:75-92) ------------------------------------------------------ */
/* Upload the reference-layout bit matrix and repack it on the device (K0) into 16-byte
 * aligned rows.  n_cols must be even (2 haplotypes per sample).                             */
STOAT_GPU_API int stoat_gpu_set_matrix(stoat_gpu_ctx* ctx, const uint8_t* bits, uint64_t n_rows, uint64_t n_cols);
/* Same, but the reference-layout bytes already live in device memory (bench "resident" leg,
 * and the upstream make_edge_matrix-on-GPU row of SURVEY §8f).                              */
STOAT_GPU_API int stoat_gpu_set_matrix_dev(stoat_gpu_ctx* ctx, const uint8_t* bits_dev, uint64_t n_rows, uint64_t n_cols);
/* binary phenotype per sample: 0 = control (g0), 1 = case (g1)
 * (parse_binary_pheno, src/arg_parser.cpp:20-95)                                            */
STOAT_GPU_API int stoat_gpu_set_binary_phenotype(stoat_gpu_ctx* ctx, const uint8_t* is_case, uint64_t n_samples);
STOAT_GPU_API int stoat_gpu_set_quantitative_phenotype(stoat_gpu_ctx* ctx, const double* y, uint64_t n_samples);
/* covariates row-major [n_samples][n_covar]; n_covar = 0 clears them.  NOTE the reference
 * indexes covariate rows by position in the FILTERED table (src/stats_test.cpp:450-452); the
 * device code reproduces that.                                                              */
STOAT_GPU_API int stoat_gpu_set_covariates(stoat_gpu_ctx* ctx, const double* covar, uint64_t n_samples, uint64_t n_covar);

/* ---- batched analyzers: host buffers in, host buffers out (the drop-in call) ----------- */
STOAT_GPU_API int stoat_gpu_binary_batch(stoat_gpu_ctx* ctx, const stoat_gpu_paths* paths,
                           const stoat_gpu_thresholds* thr, stoat_gpu_binary_out* out);
STOAT_GPU_API int stoat_gpu_quantitative_batch(stoat_gpu_ctx* ctx, const stoat_gpu_paths* paths,
                                 const stoat_gpu_thresholds* thr, stoat_gpu_quant_out* out);
/* Binary trait with covariates (logistic regression).  Needs set_binary_phenotype.  use_covariates = 0 reproduces the
 * reference as written: its covariate block declares new locals that shadow num_covariates / num_features
 * (src/stats_test.cpp:59-62), so the covariates never enter the design.  use_covariates = 1 appends the columns of
 * set_covariates to the design (by filtered row position, as the linear model does) and leaves them out of the Wald
 * tests (:148-153) — what the code evidently meant; no reference output exists for it.                           */
STOAT_GPU_API int stoat_gpu_binary_covar_batch(stoat_gpu_ctx* ctx, const stoat_gpu_paths* paths,
                                 const stoat_gpu_thresholds* thr, int use_covariates,
                                 stoat_gpu_binary_covar_out* out);
STOAT_GPU_API int stoat_gpu_binary_covar_run_staged(stoat_gpu_ctx* ctx, const stoat_gpu_thresholds* thr, int use_covariates);
STOAT_GPU_API int stoat_gpu_binary_covar_fetch(stoat_gpu_ctx* ctx, stoat_gpu_binary_covar_out* out);
STOAT_GPU_API int stoat_gpu_eqtl_batch(stoat_gpu_ctx* ctx, const stoat_gpu_paths* paths,
                         const stoat_gpu_eqtl_genes* genes, const stoat_gpu_thresholds* thr,
                         stoat_gpu_eqtl_out* out);
STOAT_GPU_API int stoat_gpu_eqtl_dense_summary(stoat_gpu_ctx* ctx, const stoat_gpu_paths* paths,
                                 const stoat_gpu_eqtl_genes* genes, const stoat_gpu_thresholds* thr,
                                 double p_threshold, stoat_gpu_eqtl_summary* out);
/* Dense scan, resident mode: the expression matrix [n_genes][n_samples] (gene major, the sampleExpresion vectors of
 * Qtl_data, src/arg_parser.hpp:49-62) is uploaded and centred once; the scan then runs over the staged paths
 * (stoat_gpu_stage_paths) without host traffic and its per-snarl summary is fetched separately.  Two-path snarls with
 * dyadic dosages and no covariates take the FP64 tensor-core contraction, every other snarl the general kernel.   */
STOAT_GPU_API int stoat_gpu_set_expression(stoat_gpu_ctx* ctx, const double* expression, uint64_t n_genes);
STOAT_GPU_API int stoat_gpu_set_expression_dev(stoat_gpu_ctx* ctx, const double* expression_dev, uint64_t n_genes);
STOAT_GPU_API int stoat_gpu_eqtl_dense_run_staged(stoat_gpu_ctx* ctx, const stoat_gpu_thresholds* thr, double p_threshold);
STOAT_GPU_API int stoat_gpu_eqtl_summary_fetch(stoat_gpu_ctx* ctx, stoat_gpu_eqtl_summary* out);

/* ---- resident mode: inputs staged once, kernels re-run without host traffic ------------ */
/* Copy the CSR to the device and keep it (and result buffers) resident in the context.     */
STOAT_GPU_API int stoat_gpu_stage_paths(stoat_gpu_ctx* ctx, const stoat_gpu_paths* paths);
/* Run K1+K2 (binary) / K3+K4 (quantitative) on the staged batch; results stay on the device */
STOAT_GPU_API int stoat_gpu_binary_run_staged(stoat_gpu_ctx* ctx, const stoat_gpu_thresholds* thr);
STOAT_GPU_API int stoat_gpu_quantitative_run_staged(stoat_gpu_ctx* ctx, const stoat_gpu_thresholds* thr);
/* Sample-sharded mode: run only the count pass, expose the device buffer that an
 * NCCL all-reduce(sum) must be applied to, then finish with the statistics pass.
 * counts layout: uint32 [2 * n_paths] = g0[0..P) then g1[0..P).                              */
STOAT_GPU_API int stoat_gpu_binary_counts_staged(stoat_gpu_ctx* ctx);
STOAT_GPU_API int stoat_gpu_binary_counts_dev(stoat_gpu_ctx* ctx, uint32_t** counts_dev, uint64_t* n_words);
STOAT_GPU_API int stoat_gpu_binary_stats_staged(stoat_gpu_ctx* ctx, const stoat_gpu_thresholds* thr);
/* Fetch the staged results into host buffers.                                               */
STOAT_GPU_API int stoat_gpu_binary_fetch(stoat_gpu_ctx* ctx, stoat_gpu_binary_out* out);
STOAT_GPU_API int stoat_gpu_quantitative_fetch(stoat_gpu_ctx* ctx, stoat_gpu_quant_out* out);

/* ---- sample-sharded quantitative run (the reference has no counterpart: its unit of
 *      parallelism is the snarl, src/snarl_analyzer.cpp:165-172) ---------------------------------
 * Every GPU holds a column range (whole samples) of the matrix and the matching slice of the
 * phenotype.  The two-path tables of quantitative_table.cpp:22-104 and the X^T X / X^T y build of
 * stats_test.cpp:444-458 are sums over samples, so the step splits around ONE all-reduce:
 *   set_quantitative_phenotype_shard   local slice + the WHOLE phenotype's mean, sum (y - mean),
 *                                      sum (y - mean)^2 and max |y - mean| (host, O(n))
 *   quantitative_moments_staged        per snarl 7 class counts (u32) + 3 sums (f64) of the local columns
 *   quantitative_partials_dev          the two device buffers to all-reduce(sum) in place (NCCL)
 *   quantitative_solve_staged          filters + OLS on the summed partials; snarls that need more
 *                                      than that come back with code STOAT_P_DEFERRED
 *   quantitative_fetch                 as usual                                                  */
STOAT_GPU_API int stoat_gpu_set_quantitative_phenotype_shard(stoat_gpu_ctx* ctx, const double* y_local, uint64_t n_local,
                                                             double mean, double sum_centred, double sum_centred_sq,
                                                             double max_abs_centred);
STOAT_GPU_API int stoat_gpu_quantitative_moments_staged(stoat_gpu_ctx* ctx, const stoat_gpu_thresholds* thr);
STOAT_GPU_API int stoat_gpu_quantitative_partials_dev(stoat_gpu_ctx* ctx, uint32_t** ints_dev, uint64_t* n_ints,
                                                      double** doubles_dev, uint64_t* n_doubles);
STOAT_GPU_API int stoat_gpu_quantitative_solve_staged(stoat_gpu_ctx* ctx, const stoat_gpu_thresholds* thr);

/* The two calls a sample-sharded run uses to finish the snarls quantitative_solve_staged deferred: every rank gathers the
 * local-column slices of the rows those snarls use (device buffer owned by the context, row i = rows[i], `pitch` bytes each),
 * the slices are exchanged with ncclAllGather and joined at their byte offsets (column ranges start at multiples of 128
 * haplotypes = 16 bytes), and the joined rows become the matrix of a second context that runs the ordinary batch call.    */
STOAT_GPU_API int stoat_gpu_gather_rows_dev(stoat_gpu_ctx* ctx, const uint32_t* rows, uint64_t n, uint8_t** out_dev,
                                            uint64_t* pitch);
/* rows already in the device layout: src_pitch bytes per row (a multiple of 16, >= n_cols / 8), bits beyond n_cols zero  */
STOAT_GPU_API int stoat_gpu_set_matrix_pitched_dev(stoat_gpu_ctx* ctx, const uint8_t* rows_dev, uint64_t n_rows,
                                                   uint64_t n_cols, uint64_t src_pitch);

/* Test hook: C [128][n_b_rows] (int32) = A [128][k] * B [n_b_rows][k]^T (int8) through the tcgen05 kernel of the dense eQTL
 * scan (n_b_rows a multiple of 256, k a multiple of 128).  Exercises the operand images, descriptors and tensor memory. */
STOAT_GPU_API int stoat_gpu_debug_umma_probe(stoat_gpu_ctx* ctx, const int8_t* a, const int8_t* b, uint32_t n_b_rows,
                                             uint32_t k, int32_t* c);

/* ---- introspection --------------------------------------------------------------------- */
/* The staged run calls (stoat_gpu_*_run_staged) only ENQUEUE work on the context's stream; this call waits for the
 * last one and reports its CUDA-event durations and counters.                                                   */
STOAT_GPU_API int stoat_gpu_last_timing(const stoat_gpu_ctx* ctx, stoat_gpu_timing* out);
/* table_ms / stats_ms averaged over the last n_runs staged runs of the same trait (at most 64 are kept); the other
 * fields as stoat_gpu_last_timing.  Lets a caller time a kernel over a whole series of back-to-back runs without a
 * host round trip between them.                                                                                 */
STOAT_GPU_API int stoat_gpu_mean_timing(stoat_gpu_ctx* ctx, uint32_t n_runs, stoat_gpu_timing* out);
/* cudaStream_t of the context as an integer (so the caller can record its own events)       */
STOAT_GPU_API uint64_t stoat_gpu_stream(const stoat_gpu_ctx* ctx);
/* The matrix from its set bits: the device half of make_edge_matrix (src/snarl_analyzer.cpp:190-260).
 * While it parses a VCF record the reference sets bit (row of the allele edge, 2 * sample + haplotype) one by one
 * (EdgeBySampleMatrix::set, src/matrix.cpp:76-84); a host that keeps those (row, column) pairs instead of the dense
 * byte matrix passes them here and the bits are set straight into the device layout (no repack pass; 8 bytes per set
 * bit go over PCIe instead of n_rows * n_cols / 8).  Pairs may repeat and come in any order.                       */
STOAT_GPU_API int stoat_gpu_set_matrix_pairs(stoat_gpu_ctx* ctx, const uint32_t* rows, const uint32_t* cols,
                                             uint64_t n_pairs, uint64_t n_rows, uint64_t n_cols);
/* Benjamini-Hochberg adjustment of a p-value column — adjust_pvalues_with_BH (src/post_processing.cpp:7-38, SURVEY.md
 * §8f-3): p_adjusted[i] in the order of p[i]; bit-identical to the reference (same two roundings per value, ties get equal
 * values).  p must be finite and >= 0: the caller drops the rows the reference's string_to_pvalue would not parse.  Host
 * buffers; n < 2^31.                                                                                                  */
STOAT_GPU_API int stoat_gpu_bh_adjust(stoat_gpu_ctx* ctx, const double* p, uint64_t n, double* p_adjusted);
/* The same adjustment WITHOUT ANY ALLOCATION.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  The scratch of the sort and the scan lives in a file-scope device array of the library (64 MiB, held by
 * every GPU that loads the library until the process exits), so n <= STOAT_GPU_BH_STATIC_MAX_N; a larger n is rejected
 * with STOAT_GPU_ERR_ARG and a message.  p and p_adjusted may be host OR device pointers (unified addressing decides);
 * the values are validated on the device (finite and >= 0, else STOAT_GPU_ERR_ARG and p_adjusted is left untouched).
 * Calls neither cudaMalloc nor cudaFree nor an allocating library routine.  The array is shared by every call and every
 * context of the process on that GPU: calls that overlap on the GPU race on it (this call synchronises its stream before
 * it returns, so calls made one after the other from one thread are safe).  Results are bit-identical to
 * stoat_gpu_bh_adjust.                                                                                               */
#define STOAT_GPU_BH_STATIC_MAX_N 1048576
STOAT_GPU_API int stoat_gpu_bh_adjust_static(stoat_gpu_ctx* ctx, const double* p, uint64_t n, double* p_adjusted);
/* bytes per repacked row on the device: ceil(n_cols / 256) * 32 (whole 32-byte sectors)         */
STOAT_GPU_API uint64_t stoat_gpu_row_pitch(const stoat_gpu_ctx* ctx);
/* read back one repacked row (tests of K0)                                                  */
STOAT_GPU_API int stoat_gpu_read_row(stoat_gpu_ctx* ctx, uint64_t row, uint8_t* out, uint64_t out_bytes);

/* ---- the value columns of the TSV rows as text ------------------------------------------------
 * The reference formats every value through an ostringstream (stoat::set_precision, src/utils.cpp:5-15) and flushes the
 * file per row (src/writer.cpp:34,49,85).  This call formats a whole column on the device: slot i = slots + 16 * i holds
 * the text of (values[i], codes[i]) — exactly what stoat_gpu_format_value returns for it — NUL padded, lengths[i] its
 * length.  codes may be NULL (all STOAT_P_VALUE).  Host buffers.                                                    */
STOAT_GPU_API int stoat_gpu_format_values(stoat_gpu_ctx* ctx, const double* values, const uint8_t* codes, uint64_t n,
                                          char* slots, uint8_t* lengths);

/* ---- host formatting helper (pure C, no device): reference strings -------------------- */
/* set_precision (src/utils.cpp:5-15) + the literal strings selected by `code`.  Returns the
 * number of characters written (excluding the NUL), or -1 if buf is too small.             */
STOAT_GPU_API int stoat_gpu_format_value(double value, uint8_t code, char* buf, size_t buf_len);

#ifdef __cplusplus
}
#endif
#endif /* STOAT_GPU_H */
