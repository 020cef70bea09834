// quant_lean.cuh — K3 for the dominant snarl shape: exactly two paths (included by kernels_quant.cu, namespace stoat).
//
// With two paths a sample's pair of haplotype counts (c0, c1), each in {0, 1, 2}, determines its table row
// (quantitative_table.cpp:71-104): row_sum = c0 + c1 and the dosages c0/row_sum, c1/row_sum.  Every class of samples is a
// bit mask built from the two presence words with a handful of LOP3s (even bit positions = samples):
//     full0   c1 == 0 < c0          d0 = 1                 full1   c0 == 0 < c1     d0 = 0
//     half    (1,1) or (2,2)        d0 = 1/2               third   (1,2)            d0 = 1/3       tthird  (2,1)  d0 = 2/3
// so X^T X and every filter sum come from POPCOUNTS of those masks (exact integers; the thirds only appear when a
// sample sits on three path-haplotypes, e.g. next to a '*' path that matches everybody), and X^T y from masked FP64 sums
// of the centred phenotype.  The last column is the baseline the reference drops (snarl_analyzer.cpp:676-684), so the
// regression has ONE dosage column (path 0) — X = [1, d0] — and K4 solves the 2 x 2 system.
//
// Exactness of the filters (snarl_analyzer.cpp:542-579): every used row sums to 1, so the table total is n_used; column
// sums are exact for dyadic dosages.  With thirds present the sums are correct to ~1e-13 relative; if a comparison is
// within 1e-9 of its threshold the snarl is handed to the general path, which replays the reference's summation order.
#pragma once

// one path's presence bits for 16-byte chunk c (code: 0xFF missing edge -> nobody, 0 no checked edge -> everybody,
// else the number of ANDed rows, <= 4)
__device__ __forceinline__ uint4 lean_path_chunk(const uint8_t* rows_s, uint32_t code, const uint32_t (&roff)[4],
                                                 uint32_t c, uint32_t n_hap) {
    if (code == 0xFFu) return make_uint4(0, 0, 0, 0);
    if (code == 0u)
        return make_uint4(valid_word(4 * c, n_hap), valid_word(4 * c + 1, n_hap), valid_word(4 * c + 2, n_hap),
                          valid_word(4 * c + 3, n_hap));
    uint4 v = *reinterpret_cast<const uint4*>(rows_s + roff[0] + c * 16);
    if (code > 1u) v = and4(v, *reinterpret_cast<const uint4*>(rows_s + roff[1] + c * 16));
    if (code > 2u) v = and4(v, *reinterpret_cast<const uint4*>(rows_s + roff[2] + c * 16));
    if (code > 3u) v = and4(v, *reinterpret_cast<const uint4*>(rows_s + roff[3] + c * 16));
    return v;
}

__device__ __forceinline__ uint32_t lean_path_word(const uint8_t* rows_s, uint32_t code, const uint32_t (&roff)[4],
                                                   uint32_t w, uint32_t n_hap) {
    if (code == 0xFFu) return 0u;
    if (code == 0u) return valid_word(w, n_hap);
    uint32_t v = *reinterpret_cast<const uint32_t*>(rows_s + roff[0] + w * 4);
    if (code > 1u) v &= *reinterpret_cast<const uint32_t*>(rows_s + roff[1] + w * 4);
    if (code > 2u) v &= *reinterpret_cast<const uint32_t*>(rows_s + roff[2] + w * 4);
    if (code > 3u) v &= *reinterpret_cast<const uint32_t*>(rows_s + roff[3] + w * 4);
    return v;
}

// masked sums of the 16 phenotype values of word w (yw = ys_s + w, stride n_words between samples)
__device__ __forceinline__ void lean_masked_sums(const double* yw, uint32_t n_words, uint32_t ma, uint32_t mb, double& sa,
                                                 double& sb) {
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        const double y = yw[i * n_words];
        if (ma & (1u << (2 * i))) sa += y;
        if (mb & (1u << (2 * i))) sb += y;
    }
}

// returns false when the snarl must take the general path
__device__ bool quant_lean2(const K3Params& p, const uint8_t* blob, const uint8_t* rows_s, const double* ys_s, uint64_t s,
                            uint32_t lane) {
    const uint32_t p0 = *reinterpret_cast<const uint32_t*>(blob + 4);
    const uint32_t code0 = blob[kBlobHeaderQ], code1 = blob[kBlobHeaderQ + 1];
    if ((code0 != 0xFFu && code0 > 4u) || (code1 != 0xFFu && code1 > 4u)) return false;
    uint32_t roff0[4], roff1[4];
    {
        uint32_t off = kBlobHeaderQ + 2;
        const uint32_t r0 = (code0 == 0xFFu) ? 0u : code0, r1 = (code1 == 0xFFu) ? 0u : code1;
#pragma unroll
        for (int r = 0; r < 4; ++r) roff0[r] = (uint32_t(r) < r0) ? uint32_t(blob[off + r]) * p.pitch : 0u;
        off += r0;
#pragma unroll
        for (int r = 0; r < 4; ++r) roff1[r] = (uint32_t(r) < r1) ? uint32_t(blob[off + r]) * p.pitch : 0u;
    }
    const uint32_t chunks = p.pitch / 16, n_words = p.pitch / 4;

    // ---- pass 1: class counts ------------------------------------------------------------------------
    uint32_t al0 = 0, al1 = 0, n_used = 0, n_full0 = 0, n_half = 0, nd = 0;
    for (uint32_t c = lane; c < chunks; c += 32) {
        const uint4 a = lean_path_chunk(rows_s, code0, roff0, c, p.n_hap);
        const uint4 b = lean_path_chunk(rows_s, code1, roff1, c, p.n_hap);
        al0 += popc4(a);
        al1 += popc4(b);
        const uint32_t av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const LeanMasks m = lean_masks(av[k], bv[k]);
            n_used += __popc(m.used);
            n_full0 += __popc(m.full0);
            n_half += __popc(m.half);
            nd |= m.third | m.tthird;
        }
    }
    uint32_t n_third = 0, n_tthird = 0;
    const bool nondyadic = __reduce_or_sync(0xffffffffu, nd) != 0u;
    if (nondyadic) {  // rare: count the two extra classes
        for (uint32_t w = lane; w < n_words; w += 32) {
            const LeanMasks m = lean_masks(lean_path_word(rows_s, code0, roff0, w, p.n_hap),
                                           lean_path_word(rows_s, code1, roff1, w, p.n_hap));
            n_third += __popc(m.third);
            n_tthird += __popc(m.tthird);
        }
        n_third = __reduce_add_sync(0xffffffffu, n_third);
        n_tthird = __reduce_add_sync(0xffffffffu, n_tthird);
    }
    if (p.n_hap < 65536u) {
        const uint32_t r0 = __reduce_add_sync(0xffffffffu, al0 | al1 << 16);
        const uint32_t r1 = __reduce_add_sync(0xffffffffu, n_used | n_full0 << 16);
        al0 = r0 & 0xFFFFu; al1 = r0 >> 16;
        n_used = r1 & 0xFFFFu; n_full0 = r1 >> 16;
    } else {
        al0 = __reduce_add_sync(0xffffffffu, al0); al1 = __reduce_add_sync(0xffffffffu, al1);
        n_used = __reduce_add_sync(0xffffffffu, n_used); n_full0 = __reduce_add_sync(0xffffffffu, n_full0);
    }
    n_half = __reduce_add_sync(0xffffffffu, n_half);
    if (lane < 2) p.out.allele_paths[p0 + lane] = lane ? al1 : al0;  // quantitative_table.cpp:54-55

    // ---- filters (snarl_analyzer.cpp:542-579) ----------------------------------------------------------
    const uint32_t n_full1 = n_used - n_full0 - n_half - n_third - n_tthird;
    const double third = 1.0 / 3.0, tthird = 2.0 / 3.0;
    const double colsum0 = double(n_full0) + 0.5 * double(n_half) + third * double(n_third) + tthird * double(n_tthird);
    const double colsum1 = double(n_full1) + 0.5 * double(n_half) + tthird * double(n_third) + third * double(n_tthird);
    bool filtered = (n_used == 0) || (al0 == 0) || (al1 == 0) || (uint64_t(n_used) < p.thr.min_individuals);
    if (!filtered) {
        const double total = double(n_used);  // every used row sums to 1
        if (nondyadic && uint64_t(n_used) == p.thr.min_haplotypes) return false;  // borderline: exact replay
        if (total < double(p.thr.min_haplotypes)) filtered = true;
        const double f0 = __ddiv_rn(colsum0, total), f1 = __ddiv_rn(colsum1, total);
        const double maf0 = fmin(f0, __dsub_rn(1.0, f0)), maf1 = fmin(f1, __dsub_rn(1.0, f1));
        if (nondyadic && (fabs(maf0 - p.thr.maf_threshold) <= 1e-9 || fabs(maf1 - p.thr.maf_threshold) <= 1e-9))
            return false;
        if (!(maf0 > p.thr.maf_threshold && maf1 > p.thr.maf_threshold)) filtered = true;
    }
    if (lane == 0) {
        p.out.filtered[s] = filtered ? 1 : 0;
        p.out.n_used[s] = n_used;
    }
    if (filtered) {
        if (lane == 0) p.mom[s].state = 0;
        return true;
    }

    // ---- pass 2: X^T y and the sums over the used samples ---------------------------------------------
    double s_full = 0.0, s_half = 0.0, s_third = 0.0, s_tthird = 0.0, su = 0.0, suu = 0.0;
    const uint32_t n_unused = p.n_samples - n_used;
    const bool walk_unused = n_unused <= n_used;
    for (uint32_t w = lane; w < n_words; w += 32) {
        const LeanMasks m = lean_masks(lean_path_word(rows_s, code0, roff0, w, p.n_hap),
                                       lean_path_word(rows_s, code1, roff1, w, p.n_hap));
        const double* yw = ys_s + w;
        lean_masked_sums(yw, n_words, m.full0, m.half, s_full, s_half);
        if (nondyadic) lean_masked_sums(yw, n_words, m.third, m.tthird, s_third, s_tthird);
        // sums over used samples: walk whichever of used / unused is the smaller set
        uint32_t walk = walk_unused ? (~m.used & 0x55555555u & valid_word(w, p.n_hap)) : m.used;
        while (walk) {
            const int b = __ffs(walk) - 1;
            walk &= walk - 1;
            const double y = yw[(b >> 1) * n_words];
            su += y;
            suu = fma(y, y, suu);
        }
    }
    double xty1 = s_full + 0.5 * s_half;
    if (nondyadic) xty1 += third * s_third + tthird * s_tthird;
    xty1 = warp_sum(xty1);
    su = warp_sum(su);
    suu = warp_sum(suu);
    if (walk_unused) {
        su = p.sy_all - su;
        suu = p.syy_all - suu;
    }
    if (lane == 0) {
        QuantMoments mom;
#pragma unroll
        for (int i = 0; i < 10; ++i) mom.xtx[i] = 0.0;
#pragma unroll
        for (int i = 0; i < 4; ++i) mom.xty[i] = 0.0;
        mom.n_used = n_used;
        mom.p = 2;
        mom.state = 1;
        mom.pad = 0;
        mom.xtx[0] = double(n_used);
        mom.xtx[1] = colsum0;
        mom.xtx[2] = double(n_full0) + 0.25 * double(n_half) + (1.0 / 9.0) * double(n_third) + (4.0 / 9.0) * double(n_tthird);
        mom.xty[0] = su;
        mom.xty[1] = xty1;
        mom.sy = su;
        mom.syy = suu;
        p.mom[s] = mom;
    }
    return true;
}
