// eqtl_dense.hpp — data layout and launchers of the dense eQTL path (kernels_eqtl.cu).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "ctx.hpp"

namespace stoat {

constexpr int kEqTileS = 128;     // snarls per CTA tile
constexpr int kEqTileG = 64;      // genes per gene tile
constexpr int kEqKT = 64;         // samples per pipeline stage (= one 16-byte chunk of a matrix row)
constexpr int kEqMaxUnused = 8;   // samples on no path a dense snarl may have (more: general path)

// Per-snarl record written by E0 (64 bytes).  state: 0 filtered, 1 dense, 2 general path.
struct __align__(16) EqtlRec {
    double inv00, inv01, inv11;  // (X^T X)^-1, X = [1, d]
    double sd, sdd;              // sum d, sum d^2 over the used samples (x00 = n_used)
    uint32_t n_used;
    uint16_t unused[kEqMaxUnused];
    uint8_t n_unused, state, pad[2];
};
static_assert(sizeof(EqtlRec) == 64, "EqtlRec must be 64 bytes");

struct EqtlDense {
    // inputs
    const double* yc;    // [n_genes][n_pad] centred expression, zero padded
    const double* tot;   // [n_genes] sum of yc over all samples
    const double* tot2;  // [n_genes] sum of yc^2
    uint32_t n_genes, n_samples, n_pad, n_kst;
    // per batch
    uint64_t n_snarls, s_pad;  // s_pad = n_snarls rounded up to kEqTileS
    uint8_t* dcode;            // [n_kst][s_pad][16]: 2-bit dosage codes of 64 samples, by SLOT (dense snarls compacted)
    EqtlRec* rec;              // [s_pad] by slot
    uint32_t* slot_snarl;      // [s_pad] slot -> snarl
    uint32_t* dense_count;     // [1] number of slots
    uint32_t* fb_list;         // snarls for the general path
    uint32_t* fb_count;        // [1]
    // gene groups (one group's rows of yc stay L2 resident while every snarl tile passes over them)
    uint32_t genes_per_group, n_groups;
    double* part_key;          // [n_groups][s_pad] best |t|
    uint32_t* part_gene;       // [n_groups][s_pad]
    uint32_t* part_cnt;        // [n_groups][s_pad]
    double p_threshold, z_threshold;  // |t| <= z_threshold can not have p < p_threshold (normal tail <= t tail)
    // outputs
    double* min_p;
    uint32_t* argmin;
    uint32_t* n_below;
};

// ---- tcgen05 int8 form (kernels_eqtl_umma.cu) ---------------------------------------------------------------------------
constexpr int kUmTileS = 128;               // slots per CTA tile = UMMA M
constexpr int kUmTileG = 32;                // genes per gene tile
constexpr int kUmTileN = kUmTileG * 8;      // limb columns per gene tile = UMMA N
constexpr int kUmKB = 128;                  // samples (bytes of a K-major int8 row) per pipeline stage
constexpr int kUmABytes = kUmTileS * kUmKB; // 16,384: image of a stage's A tile
constexpr int kUmBBytes = kUmTileN * kUmKB; // 32,768: image of a stage's B tile
struct EqtlUmma {
    uint8_t* a_img;        // [slot tiles][n_kb][16 KB]   2 d as int8, shared-memory image (K-major, 128-byte swizzle)
    const uint8_t* b_img;  // [gene tiles][n_kb][32 KB]   limbs of the centred expression, same
    const double* scale;   // [n_genes]
    uint32_t n_kb;         // n_pad / 128
    uint32_t split, rotate; // producer: 24 copies of 2 KB per stage instead of two; K blocks rotated by the slot tile
    uint32_t panel;        // slot tiles per panel of the CTA order (0 / 1: slot tile fastest over the whole grid)
    int32_t* probe_out;    // tests: raw accumulators [128][gene tiles * 32][8] of slot tile 0 instead of the statistics
};
cudaError_t launch_eqtl_umma_prepare_expression(const double* yc, uint32_t n_genes, uint32_t n_pad, double* scale,
                                                uint8_t* b_img, cudaStream_t st);
cudaError_t launch_eqtl_dense_umma(const EqtlDense& e, const EqtlUmma& u, cudaStream_t st);
size_t eqtl_umma_smem_bytes();

cudaError_t launch_eqtl_centre(const double* y, uint32_t n_genes, uint32_t n, uint32_t n_pad, double* yc, double* tot,
                               double* tot2, cudaStream_t st);
cudaError_t launch_eqtl_dosage_codes(const MatrixDev& m, const BatchDev& b, const stoat_gpu_thresholds& thr, bool allow_dense,
                                     const EqtlDense& e, QuantOut out, cudaStream_t st);
cudaError_t launch_eqtl_dense(const EqtlDense& e, int num_sms, cudaStream_t st);
cudaError_t launch_eqtl_dense_finish(const EqtlDense& e, cudaStream_t st);
size_t eqtl_dense_smem_bytes();

}  // namespace stoat
