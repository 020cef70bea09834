// dosage_masks.cuh — per-sample dosage classes of a two-path snarl as bit masks (shared by K3 dosage_moments and the
// dense eQTL path).  See quant_lean.cuh for the derivation.
#pragma once
#include <cstdint>

namespace stoat {

// bits of 32-haplotype word `w` that are real haplotypes (rows are padded to 128 bits)
__device__ __forceinline__ uint32_t valid_word(uint32_t w, uint32_t n_hap) {
    const uint32_t lo = w * 32;
    if (lo + 32 <= n_hap) return 0xFFFFFFFFu;
    if (lo >= n_hap) return 0u;
    return (1u << (n_hap - lo)) - 1u;
}

struct LeanMasks {
    uint32_t used, full0, half, third, tthird;
};

__device__ __forceinline__ LeanMasks lean_masks(uint32_t v0, uint32_t v1) {
    const uint32_t M = 0x55555555u;
    const uint32_t s0 = v0 >> 1, s1 = v1 >> 1;
    const uint32_t any0 = (v0 | s0) & M, any1 = (v1 | s1) & M;
    const uint32_t one0 = (v0 ^ s0) & M, one1 = (v1 ^ s1) & M;
    const uint32_t two0 = v0 & s0 & M, two1 = v1 & s1 & M;
    LeanMasks m;
    m.used = any0 | any1;
    m.full0 = any0 & ~any1;
    m.half = (one0 & one1) | (two0 & two1);
    m.third = one0 & two1;
    m.tthird = two0 & one1;
    return m;
}

}  // namespace stoat
