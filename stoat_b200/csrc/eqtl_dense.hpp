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

cudaError_t launch_eqtl_centre(const double* y, uint32_t n_genes, uint32_t n, uint32_t n_pad, double* yc, double* tot,
                               double* tot2, cudaStream_t st);
cudaError_t launch_eqtl_dosage_codes(const MatrixDev& m, const BatchDev& b, const stoat_gpu_thresholds& thr, bool allow_dense,
                                     const EqtlDense& e, QuantOut out, cudaStream_t st);
cudaError_t launch_eqtl_dense(const EqtlDense& e, int num_sms, cudaStream_t st);
cudaError_t launch_eqtl_dense_finish(const EqtlDense& e, cudaStream_t st);
size_t eqtl_dense_smem_bytes();

}  // namespace stoat
