// quant_general.cuh — K3G: the general quantitative / eQTL path (included by kernels_quant.cu, inside namespace stoat).
//
// One warp per snarl, any table the reference can build (<= 64 paths, <= 16 features): covariates, samples on three or
// more path-haplotypes (non-dyadic dosages), rank-deficient designs, and the eQTL gene loop.  Follows
//   process_table_quantitative / create_quantitative_table / create_eqtl_table   src/quantitative_table.cpp:22-133
//   remove_empty / filtration / combine_identical / remove_last                  src/snarl_analyzer.cpp:542-684
//   LinearRegression::linear_regression                                          src/stats_test.cpp:424-506
//   EQTLSnarlAnalyzer::analyze_and_write_snarl (table once, one OLS per gene)    src/snarl_analyzer.cpp:493-539
//
// Work layout: the presence words of every kept path are ANDed once per "superblock" (32 words = 512 samples) with
// coalesced 128-byte loads and parked in shared memory; the per-sample sweeps (lane = sample, 32 samples per block)
// then run from shared memory.  Sweep A: used rows + column sums (filters); sweep B: X^T X, X^T y, y'y; sweep C:
// residuals with the solved beta, as the reference does (:460-463).  For an eQTL snarl sweeps B (X^T y part only) and C
// repeat per gene; the table, the filters and (X^T X)^-1 are done once.
//
// Exactness of the filters: dosages count/row_sum are dyadic unless a sample sits on >= 3 path-haplotypes, and sums of
// dyadic values are exact in any order.  For non-dyadic snarls the sums are first taken in parallel order; only when a
// filter comparison is within 1e-9 of its threshold is the reference's own sequential accumulation order replayed.
#pragma once

struct GParams {
    const uint8_t* rows;
    uint32_t pitch, n_hap, n_samples;
    const double* y;      // quantitative: RAW phenotype [n] (the pseudo-inverse branch is not shift invariant);
                          // eQTL: expression [n_genes][n_samples]
    const double* covar;  // [n_samples][n_covar] row-major, indexed by FILTERED row position (stats_test.cpp:450-452)
    uint32_t n_covar;
    BatchDev b;
    const uint32_t* list;  // nullptr: every snarl
    const uint32_t* n_list_dev;
    uint64_t n_list_host;  // used when list == nullptr or n_list_dev == nullptr
    stoat_gpu_thresholds thr;
    QuantOut out;  // allele_paths [P], filtered / n_used [S]; p / r2 / beta / se / code per snarl or per (snarl, gene)
    // eQTL addressing; n_genes == 0: quantitative (one item per snarl, item index = snarl)
    uint32_t n_genes;
    const uint64_t* snarl_gene_off;  // nullptr with n_genes > 0: dense, item = s * n_genes + g
    const uint32_t* gene_idx;
    uint64_t s_begin;    // list == nullptr: snarls [s_begin, s_begin + n_list_host)
    uint64_t item_base;  // per-item outputs are written at (item - item_base): chunked dense scans
    bool items_by_slot;  // dense scan over a LIST of snarls: the items of list entry i are [i * n_genes, (i + 1) * n_genes)
    uint32_t gene_split;  // items_by_slot: a list entry is worked by gene_split warps, each a contiguous share of the genes
    // logistic regression (BinaryCovarSnarlAnalyzer, src/snarl_analyzer.cpp:392-428): the same table against the binary
    // phenotype.  is_case != nullptr selects it; the outputs are p / beta / se (r2 is left NaN).
    const uint8_t* is_case;  // [n_samples] 0 / 1
    const uint8_t* only_flagged;  // != nullptr: only the list entries e with only_flagged[e] != 0 (the rest was done elsewhere)
};

constexpr int kGWarps = 4;
constexpr int kSmallCols = 8;  // kept columns up to which sweep A keeps lane-private column sums
constexpr int kSmallP = 6;     // features up to which the moment sweeps keep lane-private accumulators (registers, no tile)
constexpr int kSmallTri = kSmallP * (kSmallP + 1) / 2;
constexpr int kSmallEnt = kSmallTri + kSmallP + 2;
struct GScratch {
    double tile[32][kMaxFeat + 1];    // X rows of the current 32 samples, last column = y
    double inv[kMaxFeat * kMaxFeat];  // (X^T X)^-1 of the snarl, column-major with leading dimension p
    double beta[kMaxFeat];
    double colsum[kMaxPathsG];
    uint32_t pw[kMaxPathsG][32];      // presence words of the kept paths for the current superblock
    uint8_t cnt[32][kMaxPathsG];      // haplotype counts of the current 32 samples per kept column
    uint8_t rs[32];                   // row sums
    uint8_t kept[kMaxPathsG];         // kept column -> path
    uint8_t fcol[kMaxPathsG];         // final column -> kept column
    uint8_t fmult[kMaxPathsG];
    uint8_t merged[kMaxPathsG];
    uint8_t same[kMaxPathsG];
    double wv[32], rv[32];            // logistic: sqrt(weight) and y - p of the current 32 samples
    uint32_t erow[kMaxPathsG][4];     // matrix rows of the kept columns' checked edges (<= 4 cached, else read through the CSR)
    uint8_t ecode[kMaxPathsG];        // 0: no checked edge (every haplotype), 1..4: cached rows, 0xFE: more than four edges
};

// word `w` (32 haplotypes) of the presence vector of path `pth` (identify_path, src/snarl_analyzer.cpp:315-356)
__device__ __forceinline__ uint32_t g_path_word(const GParams& g, uint64_t pth, uint32_t w) {
    const uint64_t a = g.b.path_edge_off[pth], z = g.b.path_edge_off[pth + 1];
    if (z == a) return valid_word(w, g.n_hap);
    uint32_t v = 0xFFFFFFFFu;
    for (uint64_t e = a; e < z; ++e) {
        const uint32_t r = g.b.edge_row[e];
        if (r == kMissingQ) return 0u;
        v &= __ldg(reinterpret_cast<const uint32_t*>(g.rows + size_t(r) * g.pitch) + w);
    }
    return v;
}

struct GSnarl {  // warp-uniform state of the snarl being processed
    uint64_t p0;
    uint32_t np, nk, n_words, n_blocks;
    uint32_t b0, b1;   // the sample blocks this warp sweeps: all of them, or its share of a team's
    uint32_t cur_sb;   // superblock whose presence words sit in sc.pw (0xFFFFFFFF: none)
};

// A TEAM of warps on one snarl (quant_general_team_kernel: the CTA is the team).  Every warp of the team runs the whole
// routine with its own scratch and takes the same decisions; only the sweeps over the samples are shared out, each warp
// taking a contiguous range of the 32-sample blocks, and every sum of a sweep is all-reduced through `buf` in a fixed
// order, so that all warps continue from bit-identical totals.  size == 1: one warp per snarl, no CTA barrier is executed.
constexpr int kGTeamWarps = 8;
struct GTeam {
    uint32_t rank, size;
    double* buf;  // [size][32]
};
__device__ __forceinline__ void team_sync(const GTeam& tm) {
    if (tm.size > 1) __syncthreads();
}
// element-wise over the lanes: lane l of every warp returns the sum over the team of its lane-l values
__device__ __forceinline__ double team_sum(const GTeam& tm, double v, uint32_t lane) {
    if (tm.size == 1) return v;
    tm.buf[tm.rank * 32 + lane] = v;
    __syncthreads();
    double r = 0.0;
    for (uint32_t t = 0; t < tm.size; ++t) r += tm.buf[t * 32 + lane];
    __syncthreads();
    return r;
}
__device__ __forceinline__ uint64_t team_and64(const GTeam& tm, uint64_t v, uint32_t lane) {
    if (tm.size == 1) return v;
    if (lane == 0) tm.buf[tm.rank * 32] = __longlong_as_double((long long)v);
    __syncthreads();
    uint64_t r = ~0ull;
    for (uint32_t t = 0; t < tm.size; ++t) r &= uint64_t(__double_as_longlong(tm.buf[t * 32]));
    __syncthreads();
    return r;
}
// the sample blocks of this warp, in order
#define G_FOR_BLOCKS(blk) for (uint32_t blk = sn.b0; blk < sn.b1; ++blk)

// word `w` of kept column c from the cached rows: the loads of a column's edges are independent (ANDed at the end)
__device__ __forceinline__ uint32_t g_cached_word(const GParams& g, const GScratch& sc, const GSnarl& sn, uint32_t c, uint32_t w) {
    const uint32_t code = sc.ecode[c];
    if (code == 0xFEu) return g_path_word(g, sn.p0 + sc.kept[c], w);
    if (code == 0u) return valid_word(w, g.n_hap);
    uint32_t v[4];
#pragma unroll
    for (int e = 0; e < 4; ++e)
        v[e] = (uint32_t(e) < code) ? __ldg(reinterpret_cast<const uint32_t*>(g.rows + size_t(sc.erow[c][e]) * g.pitch) + w) : 0xFFFFFFFFu;
    return (v[0] & v[1]) & (v[2] & v[3]);
}

// presence words of every kept column for superblock sb (lane = word).  Four columns at a time so that their row loads are
// in flight together: with the CSR walked per word (path_edge_off -> edge_row -> row, four dependent global loads per column,
// superblock and sweep) one snarl on this kernel cost ~1 ms at 5,000 samples.
__device__ __forceinline__ void g_load_superblock(const GParams& g, GScratch& sc, const GSnarl& sn, uint32_t sb,
                                                  uint32_t lane) {
    __syncwarp();
    const uint32_t w = sb * 32 + lane;
    const bool in = w < sn.n_words;
    for (uint32_t c0 = 0; c0 < sn.nk; c0 += 4) {
        uint32_t v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) v[j] = (in && c0 + j < sn.nk) ? g_cached_word(g, sc, sn, c0 + j, w) : 0u;
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (c0 + j < sn.nk) sc.pw[c0 + j][lane] = v[j];
    }
    __syncwarp();
}

// counts of the 32 samples of block blk (lane = sample) for every kept column -> sc.cnt / sc.rs; returns the row sum
__device__ __forceinline__ uint32_t g_load_block(const GParams& g, GScratch& sc, GSnarl& sn, uint32_t blk,
                                                 uint32_t lane) {
    if ((blk >> 4) != sn.cur_sb) {
        g_load_superblock(g, sc, sn, blk >> 4, lane);
        sn.cur_sb = blk >> 4;
    }
    const uint32_t i = blk * 32 + lane;
    const uint32_t wl = (blk & 15u) * 2 + (lane >> 4);
    const uint32_t sh = (lane & 15) * 2;
    uint32_t rs = 0;
    for (uint32_t c = 0; c < sn.nk; ++c) {
        const uint32_t v = sc.pw[c][wl];
        const uint32_t k = (i < g.n_samples) ? (((v >> sh) & 1u) + ((v >> (sh + 1)) & 1u)) : 0u;
        sc.cnt[lane][c] = uint8_t(k);
        rs += k;
    }
    sc.rs[lane] = uint8_t(rs > 255 ? 255 : rs);
    return rs;
}

// the same for tables of up to kSmallCols kept columns, in REGISTERS: count of column c in bits [4c, 4c + 4) of `packed`
// (a count is 0..2), no round trip through sc.cnt / sc.rs
__device__ __forceinline__ uint32_t g_load_block_small(const GParams& g, GScratch& sc, GSnarl& sn, uint32_t blk, uint32_t lane,
                                                       uint32_t& packed) {
    if ((blk >> 4) != sn.cur_sb) {
        g_load_superblock(g, sc, sn, blk >> 4, lane);
        sn.cur_sb = blk >> 4;
    }
    const uint32_t i = blk * 32 + lane;
    const uint32_t wl = (blk & 15u) * 2 + (lane >> 4);
    const uint32_t sh = (lane & 15) * 2;
    uint32_t rs = 0;
    packed = 0;
    if (i < g.n_samples) {
#pragma unroll
        for (int c = 0; c < kSmallCols; ++c)
            if (uint32_t(c) < sn.nk) {
                const uint32_t k = __popc((sc.pw[c][wl] >> sh) & 3u);
                packed |= k << (4 * c);
                rs += k;
            }
    }
    return rs;
}

// count / row_sum as the reference divides it (quantitative_table.cpp:93), without a division per column: a sample has two
// haplotypes, so a count is 0, 1 or 2, and k * rn(1 / rs) IS rn(k / rs) for those k (a product with 0, 1 or 2 is exact).
// One reciprocal per sample, then a product per column, no branches (0 for an unused sample: its counts are all 0).
__device__ __forceinline__ double g_row_rcp(uint32_t rs) {
    return __drcp_rn(double(rs ? rs : 1u));
}

// exact replay of filtration_quantitative_table's accumulation order (src/snarl_analyzer.cpp:558-576) for the rare
// non-dyadic snarl whose filter comparison is within rounding of a threshold: `total` is one running sum over
// (row, column); column sums accumulate down the rows.  Lane c owns column c (and c + 32); lane 0 owns the total.
__device__ void g_exact_sums(const GParams& g, GScratch& sc, GSnarl& sn, uint32_t lane, double& total) {
    double colsum0 = 0.0, colsum1 = 0.0, t = 0.0;
    for (uint32_t blk = 0; blk < sn.n_blocks; ++blk) {
        const uint32_t rs = g_load_block(g, sc, sn, blk, lane);
        const uint32_t used_ballot = __ballot_sync(0xffffffffu, rs > 0);
        __syncwarp();
        if (sn.nk <= uint32_t(kMaxFeat + 1)) {
            // the quotients of the block's 32 samples in parallel (lane = sample, sc.tile is free at this point), then only
            // the ADDITIONS in the reference's order: with the divisions inside the serial loop one such snarl cost ~1 ms
            for (uint32_t c = 0; c < sn.nk; ++c) {
                const uint32_t k = sc.cnt[lane][c];
                sc.tile[lane][c] = (k > 0) ? __ddiv_rn(double(k), double(rs > 255 ? 255 : rs)) : 0.0;
            }
            __syncwarp();
            if (lane < sn.nk)
                for (uint32_t i = 0; i < 32; ++i)
                    if ((used_ballot >> i) & 1u) colsum0 = __dadd_rn(colsum0, sc.tile[i][lane]);
            if (lane == 0)
                for (uint32_t i = 0; i < 32; ++i)
                    if ((used_ballot >> i) & 1u)
                        for (uint32_t c = 0; c < sn.nk; ++c) t = __dadd_rn(t, sc.tile[i][c]);
            __syncwarp();
            continue;
        }
        for (uint32_t i = 0; i < 32; ++i) {
            if (!((used_ballot >> i) & 1u)) continue;
            const double rsum = double(sc.rs[i]);
            if (lane < sn.nk) {
                const uint32_t k = sc.cnt[i][lane];
                colsum0 = __dadd_rn(colsum0, (k > 0) ? __ddiv_rn(double(k), rsum) : 0.0);
            }
            if (lane + 32 < sn.nk) {
                const uint32_t k = sc.cnt[i][lane + 32];
                colsum1 = __dadd_rn(colsum1, (k > 0) ? __ddiv_rn(double(k), rsum) : 0.0);
            }
            if (lane == 0)
                for (uint32_t c = 0; c < sn.nk; ++c) {
                    const uint32_t k = sc.cnt[i][c];
                    t = __dadd_rn(t, (k > 0) ? __ddiv_rn(double(k), rsum) : 0.0);
                }
        }
        __syncwarp();
    }
    if (lane < sn.nk) sc.colsum[lane] = colsum0;
    if (lane + 32 < sn.nk) sc.colsum[lane + 32] = colsum1;
    total = __shfl_sync(0xffffffffu, t, 0);
    __syncwarp();
}

__device__ void quant_general(const GParams& g, GScratch& sc, uint64_t s, uint64_t slot, uint32_t part, uint32_t lane,
                              const GTeam& tm) {
    GSnarl sn;
    sn.p0 = g.b.snarl_path_off[s];
    const uint64_t p1 = g.b.snarl_path_off[s + 1];
    sn.np = uint32_t(p1 - sn.p0);
    sn.n_words = g.pitch / 4;
    sn.n_blocks = (g.n_samples + 31) / 32;
    sn.nk = 0;
    sn.cur_sb = 0xFFFFFFFFu;
    {
        const uint32_t per = (sn.n_blocks + tm.size - 1) / tm.size;
        sn.b0 = tm.rank * per < sn.n_blocks ? tm.rank * per : sn.n_blocks;
        sn.b1 = sn.b0 + per < sn.n_blocks ? sn.b0 + per : sn.n_blocks;
    }
    const double nan = __longlong_as_double(0x7ff8000000000000LL);
    // items (regressions) of this snarl: one for the quantitative analyzer, one per gene of the window for eQTL
    uint64_t k0 = s, k1 = s + 1, gene_base = 0;
    if (g.n_genes) {
        const uint64_t dense_row = g.items_by_slot ? slot : s;
        k0 = g.snarl_gene_off ? g.snarl_gene_off[s] : dense_row * uint64_t(g.n_genes);
        k1 = g.snarl_gene_off ? g.snarl_gene_off[s + 1] : (dense_row + 1) * uint64_t(g.n_genes);
        if (g.items_by_slot && g.gene_split > 1) {  // this warp's share of the entry's genes
            const uint64_t share = (uint64_t(g.n_genes) + g.gene_split - 1) / g.gene_split;
            const uint64_t lo = k0 + uint64_t(part) * share;
            k1 = lo + share < k1 ? lo + share : k1;
            k0 = lo < k1 ? lo : k1;
        }
        gene_base = dense_row * uint64_t(g.n_genes);
    }
    const bool writer = tm.rank == 0;  // one warp of the team writes the results
    auto finish_none = [&](bool filtered, uint32_t n_used, uint8_t code) {
        if (lane == 0 && writer) {
            g.out.filtered[s] = filtered ? 1 : 0;
            g.out.n_used[s] = n_used;
        }
        for (uint64_t k = k0 + lane; k < k1 && writer; k += 32) {
            const uint64_t o = k - g.item_base;
            g.out.p[o] = g.out.beta[o] = g.out.se[o] = nan;
            if (g.out.r2) g.out.r2[o] = nan;
            g.out.code[o] = code;
        }
    };
    if (sn.np > uint32_t(kMaxPathsG)) {  // over the device limit: reported, never mangled
        for (uint32_t j = lane; j < sn.np && writer; j += 32) g.out.allele_paths[sn.p0 + j] = 0;
        finish_none(true, 0, kPOverLimit);
        return;
    }
    // ---- allele_paths and the kept columns (quantitative_table.cpp:37-58) --------------------------
    __syncwarp();
    const uint32_t w_first = tm.rank * 32 + lane, w_step = tm.size * 32;
    double ap0 = 0.0, ap1 = 0.0;  // lane j owns the count of path j (and j + 32)
    for (uint32_t j = 0; j < sn.np; ++j) {
        uint32_t c = 0;
        const uint64_t ea = g.b.path_edge_off[sn.p0 + j], ez = g.b.path_edge_off[sn.p0 + j + 1];
        if (ez - ea <= 4) {  // the CSR is read once per path, the row words of a pass are independent loads
            uint32_t rw[4] = {0, 0, 0, 0};
            bool missing = false;
            for (uint64_t e = ea; e < ez; ++e) {
                rw[e - ea] = g.b.edge_row[e];
                missing |= rw[e - ea] == kMissingQ;
            }
            const uint32_t ne = uint32_t(ez - ea);
            if (!missing)
                for (uint32_t w = w_first; w < sn.n_words; w += w_step) {
                    uint32_t v[4];
#pragma unroll
                    for (int e = 0; e < 4; ++e)
                        v[e] = (uint32_t(e) < ne) ? __ldg(reinterpret_cast<const uint32_t*>(g.rows + size_t(rw[e]) * g.pitch) + w)
                                                  : valid_word(w, g.n_hap);
                    c += __popc((v[0] & v[1]) & (v[2] & v[3]));
                }
        } else {
            for (uint32_t w = w_first; w < sn.n_words; w += w_step) c += __popc(g_path_word(g, sn.p0 + j, w));
        }
        c = __reduce_add_sync(0xffffffffu, c);
        if (lane == (j & 31u)) (j < 32 ? ap0 : ap1) = double(c);
    }
    ap0 = team_sum(tm, ap0, lane);
    if (sn.np > 32) ap1 = team_sum(tm, ap1, lane);
    for (uint32_t j = 0; j < sn.np; ++j) {
        const uint32_t c = uint32_t(__shfl_sync(0xffffffffu, j < 32 ? ap0 : ap1, j & 31u));
        if (lane == 0 && writer) g.out.allele_paths[sn.p0 + j] = c;
        if (c > 0) {
            if (lane == 0) sc.kept[sn.nk] = uint8_t(j);
            ++sn.nk;
        }
    }
    __syncwarp();
    const uint32_t nk = sn.nk;
    for (uint32_t c = lane; c < nk; c += 32) {  // the kept columns' edge rows, read from the CSR once
        const uint64_t pth = sn.p0 + sc.kept[c];
        const uint64_t a = g.b.path_edge_off[pth], z = g.b.path_edge_off[pth + 1];
        uint8_t code = uint8_t(z - a <= 4 ? (z - a) : 0xFEu);
        if (code != 0xFEu)
            for (uint64_t e = a; e < z; ++e) sc.erow[c][e - a] = g.b.edge_row[e];  // a kept column has no missing edge
        sc.ecode[c] = code;
    }
    __syncwarp();
    // ---- sweep A: used rows and column sums --------------------------------------------------------
    uint32_t n_used = 0;
    bool nondyadic = false;
    double colsum0 = 0.0, colsum1 = 0.0;  // lane c owns column c (and c + 32)
    double cs_lane[kSmallCols];
#pragma unroll
    for (int c = 0; c < kSmallCols; ++c) cs_lane[c] = 0.0;
    const bool fastk = nk <= uint32_t(kSmallCols);  // counts of a block in registers (g_load_block_small)
    G_FOR_BLOCKS(blk) {
        if (fastk) {  // the usual table: every lane keeps its own partial column sums, one reduction at the end
            uint32_t packed;
            const uint32_t rs = g_load_block_small(g, sc, sn, blk, lane, packed);
            n_used += __popc(__ballot_sync(0xffffffffu, rs > 0));
            nondyadic |= __any_sync(0xffffffffu, (rs & (rs - 1)) != 0);
            const double rinv = g_row_rcp(rs);
#pragma unroll
            for (int c = 0; c < kSmallCols; ++c)  // (columns >= nk hold 0; the product is exact, so the fma adds what the reference adds)
                cs_lane[c] = fma(double((packed >> (4 * c)) & 15u), rinv, cs_lane[c]);
            continue;
        }
        const uint32_t rs = g_load_block(g, sc, sn, blk, lane);
        const uint32_t used_ballot = __ballot_sync(0xffffffffu, rs > 0);
        n_used += __popc(used_ballot);
        nondyadic |= __any_sync(0xffffffffu, rs >= 3 && (rs & (rs - 1)) != 0);
        const double rsum = double(rs);
        for (uint32_t c = 0; c < nk; ++c) {
            const uint32_t k = sc.cnt[lane][c];
            const double d = (k > 0) ? __ddiv_rn(double(k), rsum) : 0.0;  // quantitative_table.cpp:93
            const double sum = warp_sum(d);
            if (lane == (c & 31u)) {
                if (c < 32) colsum0 += sum;
                else colsum1 += sum;
            }
        }
    }
    if (nk <= uint32_t(kSmallCols)) {
#pragma unroll
        for (int c = 0; c < kSmallCols; ++c) {
            const double tot = warp_sum(cs_lane[c]);  // dyadic terms: exact in any order (else the replay below decides)
            if (lane == uint32_t(c)) colsum0 = tot;
        }
    }
    uint32_t rank0 = 0;  // used samples in the blocks before this warp's: covariate rows go by FILTERED position
    if (tm.size > 1) {
        colsum0 = team_sum(tm, colsum0, lane);
        if (nk > 32) colsum1 = team_sum(tm, colsum1, lane);
        tm.buf[tm.rank * 32 + lane] = lane == 0 ? double(n_used) : (lane == 1 && nondyadic ? 1.0 : 0.0);
        __syncthreads();
        double tot_used = 0.0, tot_nd = 0.0;
        for (uint32_t t = 0; t < tm.size; ++t) {
            if (t == tm.rank) rank0 = uint32_t(tot_used);
            tot_used += tm.buf[t * 32];
            tot_nd += tm.buf[t * 32 + 1];
        }
        __syncthreads();
        n_used = uint32_t(tot_used);
        nondyadic = tot_nd > 0.0;
    }
    // remove_empty_columns is a no-op here (empty paths were never added); filtration (:549)
    if (n_used == 0 || nk < 2 || uint64_t(n_used) < g.thr.min_individuals) {
        finish_none(true, n_used, kPNa);
        return;
    }
    __syncwarp();
    if (lane < nk) sc.colsum[lane] = colsum0;
    if (lane + 32 < nk) sc.colsum[lane + 32] = colsum1;
    __syncwarp();
    double total = 0.0;
    for (uint32_t c = 0; c < nk; ++c) total += sc.colsum[c];  // exact when every term is dyadic
    auto run_filters = [&](double tot, bool& borderline) -> bool {
        borderline = false;
        const double eps = 1e-9;
        const double mh = double(g.thr.min_haplotypes);
        if (fabs(tot - mh) <= eps * fmax(tot, 1.0)) borderline = true;
        bool filt = tot < mh;  // :565
        int above = 0;
        for (uint32_t c = 0; c < nk; ++c) {  // :570-576
            const double freq = __ddiv_rn(sc.colsum[c], tot);
            const double maf = fmin(freq, __dsub_rn(1.0, freq));
            if (fabs(maf - g.thr.maf_threshold) <= eps) borderline = true;
            if (maf > g.thr.maf_threshold) ++above;
        }
        return filt || above < 2;
    };
    bool borderline = false;
    bool filtered = run_filters(total, borderline);
    if (nondyadic && borderline) {
        g_exact_sums(g, sc, sn, lane, total);
        filtered = run_filters(total, borderline);
    }
    if (filtered) {
        finish_none(true, n_used, kPNa);
        return;
    }
    // ---- combine identical columns (:620-674): lane <-> candidate column b > a ----------------------
    uint32_t nf = 0;
    if (lane == 0)
        for (uint32_t c = 0; c < nk; ++c) sc.merged[c] = 0;
    __syncwarp();
    for (uint32_t a = 0; a < nk; ++a) {
        if (sc.merged[a]) continue;
        uint32_t mult = 1;
        if (nk >= 3 && a + 1 < nk) {
            // candidates b > a that may equal column a; lane = sample, one vote per candidate and block
            uint64_t cand = 0;
            for (uint32_t b2 = a + 1; b2 < nk; ++b2)
                if (!sc.merged[b2]) cand |= 1ull << b2;
            for (uint32_t blk = sn.b0; blk < sn.b1 && cand; ++blk) {
                g_load_block(g, sc, sn, blk, lane);
                __syncwarp();
                const uint8_t ka = sc.cnt[lane][a];
                uint64_t c2 = cand;
                while (c2) {
                    const int b2 = __ffsll((long long)c2) - 1;
                    c2 &= c2 - 1;
                    if (__any_sync(0xffffffffu, sc.cnt[lane][b2] != ka)) cand &= ~(1ull << b2);
                }
                __syncwarp();
            }
            cand = team_and64(tm, cand, lane);
            if (lane == 0)
                for (uint32_t b2 = a + 1; b2 < nk; ++b2)
                    if ((cand >> b2) & 1ull) sc.merged[b2] = 1;
            mult += uint32_t(__popcll(cand));
            __syncwarp();
        }
        if (lane == 0) {
            sc.fcol[nf] = uint8_t(a);
            sc.fmult[nf] = uint8_t(mult);
        }
        ++nf;
        __syncwarp();
    }
    const uint32_t k_reg = nf - 1;  // remove_last_columns (:676-684)
    const uint32_t p = 1 + k_reg + g.n_covar;
    if (k_reg == 0) {  // the reference indexes past the end here (stats_test.cpp:487-489)
        finish_none(false, n_used, kPUndef);
        return;
    }
    if (p > uint32_t(kMaxFeat)) {
        finish_none(false, n_used, kPOverLimit);
        return;
    }
    if (lane == 0 && writer) {
        g.out.filtered[s] = 0;
        g.out.n_used[s] = n_used;
    }
    // ---- regressions ---------------------------------------------------------------------------------
    // entries e of [packed lower triangle of X^T X | X^T y | y'y | sum y] are owned by lanes (e % 32)
    const uint32_t n_tri = p * (p + 1) / 2;
    const uint32_t n_ent = n_tri + p + 2;
    const bool small_p = p <= uint32_t(kSmallP);
    uint32_t er[6], ec[6];  // ceil((136 + 16 + 2) / 32) = 5
#pragma unroll
    for (int q = 0; q < 6; ++q) {
        const uint32_t e = lane + 32u * q;
        uint32_t r = 0;
        while (!small_p && (r + 1) * (r + 2) / 2 <= e) ++r;  // (the register sweeps of small designs do not use them)
        er[q] = r;
        ec[q] = e - r * (r + 1) / 2;
    }
    uint32_t rank_base = rank0;
    auto build_tile = [&](uint32_t blk, uint32_t used_ballot, const double* y) {
        // lane = sample: write its X row (zeros when unused)
        const bool used = (used_ballot >> lane) & 1u;
        const uint32_t i = blk * 32 + lane;
        const uint32_t rank = rank_base + __popc(used_ballot & ((1u << lane) - 1u));
        const double rsum = double(sc.rs[lane]);
        sc.tile[lane][0] = used ? 1.0 : 0.0;
        for (uint32_t a = 0; a < k_reg; ++a) {
            const uint32_t k = sc.cnt[lane][sc.fcol[a]];
            // merged columns are sums of identical columns (:653-658): mult * (k / row_sum)
            const double d = (used && k > 0) ? __ddiv_rn(double(k), rsum) : 0.0;
            double v = d;
            for (uint32_t m = 1; m < sc.fmult[a]; ++m) v = __dadd_rn(v, d);
            sc.tile[lane][1 + a] = v;
        }
        for (uint32_t c = 0; c < g.n_covar; ++c)
            // covariate rows are taken by position in the FILTERED table (stats_test.cpp:450-452)
            sc.tile[lane][1 + k_reg + c] = used ? g.covar[size_t(rank) * g.n_covar + c] : 0.0;
        sc.tile[lane][kMaxFeat] = (used && i < g.n_samples) ? (g.is_case ? (g.is_case[i] ? 1.0 : 0.0) : y[i]) : 0.0;
    };
    // the same row in REGISTERS (lane = sample), for designs of up to kSmallP features: the moment sweeps then keep
    // lane-private accumulators and reduce once per sweep, instead of 32-step loops over a shared-memory tile per block
    // column maps of the small design in registers; `fastk`: the block's counts come packed in a register as well
    uint32_t fc[kSmallP - 1], fm[kSmallP - 1], fmask[kSmallP - 1];
    bool has_mult = false;
#pragma unroll
    for (int a = 0; a < kSmallP - 1; ++a) {
        const bool on = small_p && uint32_t(a) < k_reg;
        fc[a] = on ? sc.fcol[a] : 0u;
        fm[a] = on ? sc.fmult[a] : 1u;
        fmask[a] = on ? 15u : 0u;  // columns the design does not have read as count 0
        has_mult |= fm[a] > 1;
    }
    // value of sample blk * 32 + lane: the phenotype, or the case flag of the logistic mode
    auto load_y = [&](uint32_t blk, const double* y) __attribute__((always_inline)) -> double {
        const uint32_t i = blk * 32 + lane;
        if (blk >= sn.b1 || i >= g.n_samples) return 0.0;
        return g.is_case ? (g.is_case[i] ? 1.0 : 0.0) : y[i];
    };
    // one block of the small design: counts (registers when fastk), the row in registers, y from the value loaded one
    // block ahead (`y_next` in, refilled for blk + 1 before the arithmetic of this block so that the load is hidden)
    auto build_row = [&](uint32_t blk, const double* y, double& y_next, double (&x)[kSmallP], double& yv)
                         __attribute__((always_inline)) -> bool {
        uint32_t packed = 0;
        const uint32_t rs = fastk ? g_load_block_small(g, sc, sn, blk, lane, packed) : g_load_block(g, sc, sn, blk, lane);
        const double y_cur = y_next;
        y_next = load_y(blk + 1, y);
        const uint32_t used_ballot = __ballot_sync(0xffffffffu, rs > 0);
        const bool used = rs > 0;
        const uint32_t rank = rank_base + __popc(used_ballot & ((1u << lane) - 1u));
        rank_base += __popc(used_ballot);
        const double rinv = g_row_rcp(rs > 255 ? 255 : rs);  // (the clamp as sc.rs holds it)
        x[0] = used ? 1.0 : 0.0;
#pragma unroll
        for (int a = 0; a < kSmallP - 1; ++a) {
            const uint32_t k = (fastk ? (packed >> (4 * fc[a])) : uint32_t(sc.cnt[lane][fc[a]])) & fmask[a];
            x[1 + a] = double(k) * rinv;  // k / row_sum (0 for an unused sample)
        }
        if (has_mult) {  // merged columns are sums of identical columns (:653-658): d + d + ... as the reference adds them
#pragma unroll
            for (int a = 0; a < kSmallP - 1; ++a) {
                const double d = x[1 + a];
                for (uint32_t m = 1; m < fm[a]; ++m) x[1 + a] = __dadd_rn(x[1 + a], d);
            }
        }
#pragma unroll
        for (int c = 0; c < kSmallP - 2; ++c)
            if (uint32_t(c) < g.n_covar && 1 + k_reg + uint32_t(c) < uint32_t(kSmallP)) {
                const double cv = used ? g.covar[size_t(rank) * g.n_covar + c] : 0.0;  // by FILTERED row position (:450-452)
#pragma unroll
                for (int r = 2; r < kSmallP; ++r)
                    if (uint32_t(r) == 1 + k_reg + uint32_t(c)) x[r] = cv;
            }
        yv = used ? y_cur : 0.0;
        return used;
    };
    if (g.is_case) {
        // ---- LogisticRegression::logistic_regression (src/stats_test.cpp:49-176) ----------------------
        // IRLS: every pass sweeps the samples once for the Hessian X^T W X + l2 I and the gradient X^T (y - p) - l2 beta
        // at the current beta (lanes own the entries, as in sweep B); lane 0 then takes the Newton step (pivoted LDLT,
        // :107-111).  After the step that meets the tolerance (:113) one more pass gives the Hessian whose inverse is the
        // covariance (:124-139).
        constexpr double kTol = 1e-6, kL2 = 1e-4, kEps = 1e-8;  // src/stats_test.hpp:109-112
        constexpr int kMaxIter = 100;
        const uint32_t n_tri = p * (p + 1) / 2;
        const uint32_t n_ent = n_tri + p;
        double* beta_old = sc.inv;  // free here: no (X^T X)^-1 in this mode
        if (lane < p) {
            sc.beta[lane] = 0.0;
            beta_old[lane] = 0.0;
        }
        __syncwarp();
        int steps = 0;
        int state = 0;  // 0: iterate, 1: converged (this pass is the covariance pass), 2: done
        while (state != 2) {
            double* flat = &sc.tile[0][0];
            rank_base = rank0;
            if (small_p) {
                double acc[kSmallTri + kSmallP];
#pragma unroll
                for (int e = 0; e < kSmallTri + kSmallP; ++e) acc[e] = 0.0;
                double bta[kSmallP];
#pragma unroll
                for (int r = 0; r < kSmallP; ++r) bta[r] = (uint32_t(r) < p) ? sc.beta[r] : 0.0;
                double y_next = load_y(sn.b0, nullptr);
                G_FOR_BLOCKS(blk) {
                    double x[kSmallP], yv;
                    const bool used = build_row(blk, nullptr, y_next, x, yv);
                    double z = 0.0;
#pragma unroll
                    for (int r = 0; r < kSmallP; ++r) z = fma(x[r], bta[r], z);
                    const double pr = 1.0 / (1.0 + exp(-z));                 // :27-29
                    const double w = fmax(kEps, fmin(1.0, pr * (1.0 - pr)));  // :93
                    const double sw = used ? sqrt(w) : 0.0;                   // X_weighted.row(i) *= sqrt(w), :97-98
                    const double rv = used ? yv - pr : 0.0;
                    int e = 0;
#pragma unroll
                    for (int r = 0; r < kSmallP; ++r) {
#pragma unroll
                        for (int c = 0; c <= r; ++c, ++e) acc[e] = fma(x[r] * sw, x[c] * sw, acc[e]);
                        acc[kSmallTri + r] = fma(x[r], rv, acc[kSmallTri + r]);
                    }
                }
#pragma unroll
                double mine = 0.0;  // lane e takes entry e (27 of them), one team reduction for all
#pragma unroll
                for (int e = 0; e < kSmallTri + kSmallP; ++e) {
                    const double v = warp_sum(acc[e]);
                    if (lane == uint32_t(e)) mine = v;
                }
                mine = team_sum(tm, mine, lane);
                if (lane < uint32_t(kSmallTri)) {
                    if (lane < n_tri) flat[lane] = mine;
                } else if (lane < uint32_t(kSmallTri + kSmallP) && lane - kSmallTri < p) {
                    flat[n_tri + (lane - kSmallTri)] = mine;
                }
                __syncwarp();
            } else {
            double accv[6] = {0, 0, 0, 0, 0, 0};
            G_FOR_BLOCKS(blk) {
                const uint32_t rs = g_load_block(g, sc, sn, blk, lane);
                const uint32_t used_ballot = __ballot_sync(0xffffffffu, rs > 0);
                __syncwarp();
                build_tile(blk, used_ballot, nullptr);
                rank_base += __popc(used_ballot);
                {
                    double z = 0.0;
                    for (uint32_t r = 0; r < p; ++r) z = fma(sc.tile[lane][r], sc.beta[r], z);
                    const double pr = 1.0 / (1.0 + exp(-z));                 // :27-29
                    const double w = fmax(kEps, fmin(1.0, pr * (1.0 - pr)));  // :93
                    const bool used = (used_ballot >> lane) & 1u;
                    sc.wv[lane] = used ? sqrt(w) : 0.0;                       // X_weighted.row(i) *= sqrt(w), :97-98
                    sc.rv[lane] = used ? sc.tile[lane][kMaxFeat] - pr : 0.0;
                }
                __syncwarp();
#pragma unroll
                for (int q = 0; q < 6; ++q) {
                    const uint32_t e = lane + 32u * q;
                    if (e >= n_ent) continue;
                    double a2 = 0.0;
                    if (e < n_tri) {
                        for (uint32_t i = 0; i < 32; ++i)
                            a2 = fma(sc.tile[i][er[q]] * sc.wv[i], sc.tile[i][ec[q]] * sc.wv[i], a2);
                    } else {
                        const uint32_t r = e - n_tri;
                        for (uint32_t i = 0; i < 32; ++i) a2 = fma(sc.tile[i][r], sc.rv[i], a2);
                    }
                    accv[q] += a2;
                }
                __syncwarp();
            }
#pragma unroll
            for (int q = 0; q < 6; ++q)
                if (32u * q < n_ent) {  // warp-uniform condition: the team reduction holds a barrier
                    const double v = team_sum(tm, accv[q], lane);
                    if (lane + 32u * q < n_ent) flat[lane + 32u * q] = v;
                }
            __syncwarp();
            }
            if (lane == 0) {
                SmallMat<kMaxFeat> Hm;
                uint32_t e = 0;
                for (uint32_t r = 0; r < p; ++r)
                    for (uint32_t c = 0; c <= r; ++c, ++e) {
                        Hm(r, c) = flat[e];
                        Hm(c, r) = flat[e];
                    }
                for (uint32_t r = 0; r < p; ++r) Hm(r, r) += kL2;  // :101
                uint8_t code = kPValue;
                double out_p = nan, out_b = nan, out_se = nan;
                if (state == 1) {
                    double diag[kMaxFeat];
                    lu_inverse_diag<kMaxFeat>(Hm, int(p), diag);  // hessian.inverse(), :138
                    double pv[kMaxFeat];
                    const uint32_t n_tests = k_reg;  // covariate columns are not tested (:150)
                    for (uint32_t i = 0; i < n_tests; ++i) {
                        const double zs = sc.beta[1 + i] / sqrt(diag[1 + i]);
                        pv[i] = 2.0 * (1.0 - normal_cdf_like_reference(fabs(zs)));  // :151-152
                    }
                    uint32_t best = 0;
                    out_p = pv[0];
                    if (n_tests > 1) {  // adjusted_holm (src/utils.cpp:61-87) + the smallest adjusted value (:169-175)
                        uint8_t ord[kMaxFeat];
                        for (uint32_t i = 0; i < n_tests; ++i) {
                            uint32_t j = i;
                            while (j > 0 && pv[ord[j - 1]] > pv[i]) {
                                ord[j] = ord[j - 1];
                                --j;
                            }
                            ord[j] = uint8_t(i);
                        }
                        double prev = 0.0, best_v = 0.0;
                        bool have = false;
                        for (uint32_t i = 0; i < n_tests; ++i) {
                            const double raw = fmin(double(n_tests - i) * pv[ord[i]], 1.0);
                            const double adj = fmax(prev, raw);
                            prev = adj;
                            // std::min_element over the ORIGINAL order: the first index holding the minimum
                            if (!have || adj < best_v || (adj == best_v && ord[i] < best)) {
                                best_v = adj;
                                best = ord[i];
                                have = true;
                            }
                        }
                        out_p = best_v;
                    }
                    out_b = sc.beta[1 + best];
                    out_se = sqrt(diag[1 + best]);
                    state = 2;
                } else {
                    double gvec[kMaxFeat];
                    for (uint32_t r = 0; r < p; ++r) gvec[r] = flat[n_tri + r] - kL2 * sc.beta[r];  // :105
                    int tr[kMaxFeat];
                    ldlt_factor<kMaxFeat>(Hm, int(p), tr);
                    bool bad = false;
                    for (uint32_t r = 0; r < p; ++r) bad |= isnan(Hm(r, r));
                    if (bad) {
                        code = kPNa;  // ldlt.info() != Success, :107
                        state = 2;
                    } else {
                        ldlt_solve<kMaxFeat>(Hm, int(p), tr, gvec);
                        double nrm = 0.0;
                        for (uint32_t r = 0; r < p; ++r) {
                            sc.beta[r] += gvec[r];
                            const double d = sc.beta[r] - beta_old[r];
                            nrm = fma(d, d, nrm);
                        }
                        ++steps;
                        if (sqrt(nrm) < kTol) {
                            state = 1;
                        } else if (steps >= kMaxIter) {
                            code = kPNa;  // not converged, :120
                            state = 2;
                        } else {
                            for (uint32_t r = 0; r < p; ++r) beta_old[r] = sc.beta[r];
                        }
                    }
                }
                if (state == 2 && writer) {
                    const uint64_t oi = s - g.item_base;
                    g.out.p[oi] = out_p;
                    if (g.out.r2) g.out.r2[oi] = nan;
                    g.out.beta[oi] = out_b;
                    g.out.se[oi] = out_se;
                    g.out.code[oi] = code;
                }
            }
            state = __shfl_sync(0xffffffffu, state, 0);
            steps = __shfl_sync(0xffffffffu, steps, 0);
            __syncwarp();
        }
        return;
    }
    for (uint64_t item = k0; item < k1; ++item) {
        const bool first = (item == k0);
        const double* y = g.y;
        if (g.n_genes) {
            const uint64_t gene = g.gene_idx ? uint64_t(g.gene_idx[item]) : (item - gene_base);
            y = g.y + gene * uint64_t(g.n_samples);  // retain_indices(expression, index_filtered), :510-511
        }
        // ---- sweep B: X^T X (first item only), X^T y, y'y, sum y ------------------------------------
        double* flat = &sc.tile[0][0];
        rank_base = rank0;
        if (small_p) {
            double acc[kSmallEnt];
#pragma unroll
            for (int e = 0; e < kSmallEnt; ++e) acc[e] = 0.0;
            double y_next = load_y(sn.b0, y);
            G_FOR_BLOCKS(blk) {
                double x[kSmallP], yv;
                build_row(blk, y, y_next, x, yv);
                int e = 0;
#pragma unroll
                for (int r = 0; r < kSmallP; ++r) {
                    if (first) {
#pragma unroll
                        for (int c = 0; c <= r; ++c) acc[e + c] = fma(x[r], x[c], acc[e + c]);
                    }
                    e += r + 1;
                    acc[kSmallTri + r] = fma(x[r], yv, acc[kSmallTri + r]);
                }
                acc[kSmallTri + kSmallP] = fma(yv, yv, acc[kSmallTri + kSmallP]);
                acc[kSmallTri + kSmallP + 1] += yv;
            }
#pragma unroll
            double mine = 0.0;  // lane e takes entry e (29 of them), one team reduction for all
#pragma unroll
            for (int e = 0; e < kSmallEnt; ++e) {
                if (e < kSmallTri && !first) continue;
                const double v = warp_sum(acc[e]);
                if (lane == uint32_t(e)) mine = v;
            }
            mine = team_sum(tm, mine, lane);
            if (lane < uint32_t(kSmallTri)) {
                if (first && lane < n_tri) flat[lane] = mine;
            } else if (lane < uint32_t(kSmallTri + kSmallP)) {
                if (lane - kSmallTri < p) flat[n_tri + (lane - kSmallTri)] = mine;
            } else if (lane < uint32_t(kSmallEnt)) {
                flat[n_tri + p + (lane - kSmallTri - kSmallP)] = mine;
            }
            __syncwarp();
        } else {
        double accv[6] = {0, 0, 0, 0, 0, 0};
        G_FOR_BLOCKS(blk) {
            const uint32_t rs = g_load_block(g, sc, sn, blk, lane);
            const uint32_t used_ballot = __ballot_sync(0xffffffffu, rs > 0);
            __syncwarp();
            build_tile(blk, used_ballot, y);
            rank_base += __popc(used_ballot);
            __syncwarp();
#pragma unroll
            for (int q = 0; q < 6; ++q) {
                const uint32_t e = lane + 32u * q;
                if (e >= n_ent || (!first && e < n_tri)) continue;
                double a2 = 0.0;
                if (e < n_tri) {
                    for (uint32_t i = 0; i < 32; ++i) a2 = fma(sc.tile[i][er[q]], sc.tile[i][ec[q]], a2);
                } else if (e < n_tri + p) {
                    const uint32_t r = e - n_tri;
                    for (uint32_t i = 0; i < 32; ++i) a2 = fma(sc.tile[i][r], sc.tile[i][kMaxFeat], a2);
                } else if (e == n_tri + p) {
                    for (uint32_t i = 0; i < 32; ++i) a2 = fma(sc.tile[i][kMaxFeat], sc.tile[i][kMaxFeat], a2);
                } else {
                    for (uint32_t i = 0; i < 32; ++i) a2 += sc.tile[i][kMaxFeat];
                }
                accv[q] += a2;
            }
            __syncwarp();
        }
        // ---- solve on lane 0: gather the entries through shared memory (reuse the tile) --------------
#pragma unroll
        for (int q = 0; q < 6; ++q)
            if (32u * q < n_ent) {  // warp-uniform condition: the team reduction holds a barrier
                const double v = team_sum(tm, accv[q], lane);
                if (lane + 32u * q < n_ent && (first || lane + 32u * q >= n_tri)) flat[lane + 32u * q] = v;
            }
        __syncwarp();
        }
        double sum_y = 0.0;
        if (lane == 0) {
            if (first) {
                SmallMat<kMaxFeat> X, inv;
                uint32_t e = 0;
                for (uint32_t r = 0; r < p; ++r)
                    for (uint32_t c = 0; c <= r; ++c, ++e) {
                        X(r, c) = flat[e];
                        X(c, r) = flat[e];
                    }
                xtx_inverse<kMaxFeat>(X, int(p), inv);
                for (uint32_t r = 0; r < p; ++r)
                    for (uint32_t c = 0; c < p; ++c) sc.inv[r + c * p] = inv(r, c);
            }
            sum_y = flat[n_tri + p + 1];
            for (uint32_t r = 0; r < p; ++r) {
                double b2 = 0.0;
                for (uint32_t c = 0; c < p; ++c) b2 += sc.inv[r + c * p] * flat[n_tri + c];
                sc.beta[r] = b2;
            }
        }
        __syncwarp();
        sum_y = __shfl_sync(0xffffffffu, sum_y, 0);
        const double y_mean = sum_y / double(n_used);
        // ---- sweep C: residuals (stats_test.cpp:459-465) ----------------------------------------------
        double rss = 0.0, tss = 0.0;
        rank_base = rank0;
        if (small_p) {
            double bta[kSmallP];
#pragma unroll
            for (int r = 0; r < kSmallP; ++r) bta[r] = (uint32_t(r) < p) ? sc.beta[r] : 0.0;
            double y_next = load_y(sn.b0, y);
            G_FOR_BLOCKS(blk) {
                double x[kSmallP], yv;
                const bool used = build_row(blk, y, y_next, x, yv);
                if (used) {
                    double pred = 0.0;
#pragma unroll
                    for (int r = 0; r < kSmallP; ++r) pred = fma(x[r], bta[r], pred);
                    const double res = yv - pred;
                    rss = fma(res, res, rss);
                    const double dm = yv - y_mean;
                    tss = fma(dm, dm, tss);
                }
            }
        } else {
        G_FOR_BLOCKS(blk) {
            const uint32_t rs = g_load_block(g, sc, sn, blk, lane);
            const uint32_t used_ballot = __ballot_sync(0xffffffffu, rs > 0);
            __syncwarp();
            build_tile(blk, used_ballot, y);
            rank_base += __popc(used_ballot);
            __syncwarp();
            if ((used_ballot >> lane) & 1u) {
                double pred = 0.0;
                for (uint32_t r = 0; r < p; ++r) pred = fma(sc.tile[lane][r], sc.beta[r], pred);
                const double yv = sc.tile[lane][kMaxFeat];
                const double res = yv - pred;
                rss = fma(res, res, rss);
                const double dm = yv - y_mean;
                tss = fma(dm, dm, tss);
            }
            __syncwarp();
        }
        }
        rss = warp_sum(rss);
        tss = warp_sum(tss);
        if (tm.size > 1) {
            const double aux = team_sum(tm, lane == 0 ? rss : (lane == 1 ? tss : 0.0), lane);
            rss = __shfl_sync(0xffffffffu, aux, 0);
            tss = __shfl_sync(0xffffffffu, aux, 1);
        }
        if (lane == 0 && writer) {
            const OlsOut o = ols_statistics(sc.beta[1], sc.inv[1 + p], rss, tss, n_used, int(p));
            const uint64_t oi = item - g.item_base;
            g.out.p[oi] = o.p;
            g.out.r2[oi] = o.r2;
            g.out.beta[oi] = o.beta;
            g.out.se[oi] = o.se;
            g.out.code[oi] = kPValue;
        }
        __syncwarp();
    }
}

// n > kGTeamWarps-wide teams pay off only while there are fewer snarls than the grid has warp slots
__device__ __forceinline__ bool g_team_mode(const GParams& g, uint64_t n, uint32_t team_limit) {
    return team_limit > 0 && n <= team_limit;
}

__global__ void __launch_bounds__(kGWarps * 32) quant_general_kernel(const GParams g, uint32_t team_limit) {
    extern __shared__ __align__(16) uint8_t g_smem[];
    GScratch* scratch = reinterpret_cast<GScratch*>(g_smem);
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint64_t gw = uint64_t(blockIdx.x) * kGWarps + warp;
    const uint64_t n_warps = uint64_t(gridDim.x) * kGWarps;
    const uint64_t split = (g.items_by_slot && g.gene_split > 1) ? g.gene_split : 1;
    const uint64_t n_list = (g.list && g.n_list_dev) ? uint64_t(*g.n_list_dev) : g.n_list_host;
    if (g_team_mode(g, n_list, team_limit)) return;  // quant_general_team_kernel has them
    const uint64_t n = n_list * split;
    const GTeam tm{0u, 1u, nullptr};
    for (uint64_t i = gw; i < n; i += n_warps) {
        const uint64_t entry = i / split;
        if (g.only_flagged && !g.only_flagged[entry]) continue;
        const uint64_t s = g.list ? g.list[entry] : g.s_begin + entry;
        quant_general(g, scratch[warp], s, entry, uint32_t(i - entry * split), lane, tm);
    }
}

// The short list (a handful of stray snarls behind the integer-MMA kernels, a small batch): one CTA of kGTeamWarps warps
// per snarl.  A lone warp is latency bound (~0.7 ms for one snarl at 5,000 samples, 3.4 ms at 25,000), so the step
// waited on a single warp; here the sample sweeps of the snarl are spread over the team.
__global__ void __launch_bounds__(kGTeamWarps * 32) quant_general_team_kernel(const GParams g, uint32_t team_limit) {
    extern __shared__ __align__(16) uint8_t g_smem[];
    GScratch* scratch = reinterpret_cast<GScratch*>(g_smem);
    double* buf = reinterpret_cast<double*>(g_smem + sizeof(GScratch) * kGTeamWarps);
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint64_t n_list = (g.list && g.n_list_dev) ? uint64_t(*g.n_list_dev) : g.n_list_host;
    if (!g_team_mode(g, n_list, team_limit)) return;
    const GTeam tm{warp, uint32_t(kGTeamWarps), buf};
    for (uint64_t entry = blockIdx.x; entry < n_list; entry += gridDim.x) {
        if (g.only_flagged && !g.only_flagged[entry]) continue;  // (CTA-uniform)
        const uint64_t s = g.list ? g.list[entry] : g.s_begin + entry;
        quant_general(g, scratch[warp], s, entry, 0u, lane, tm);
        __syncthreads();
    }
}

inline cudaError_t launch_quant_general(const GParams& g, uint64_t max_items, int num_sms, cudaStream_t st) {
    const int smem = int(sizeof(GScratch) * kGWarps);
    cudaError_t e = cudaFuncSetAttribute(quant_general_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
    // teams while they finish sooner than one warp per snarl would: ceil(n / SMs) / 8 against ceil(n / (12 SMs))
    static const bool no_team = [] {
        const char* v = getenv("STOAT_GPU_K3G_TEAM");
        return v && v[0] == '0';
    }();
    const bool team_ok = !no_team && !(g.items_by_slot && g.gene_split > 1);
    const uint32_t team_limit = team_ok ? uint32_t(num_sms) * 6u : 0u;
    const bool count_on_device = g.list && g.n_list_dev;
    if (team_ok && (count_on_device || g.n_list_host <= team_limit)) {
        const int tsmem = int(sizeof(GScratch) * kGTeamWarps + sizeof(double) * kGTeamWarps * 32);
        e = cudaFuncSetAttribute(quant_general_team_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, tsmem);
        if (e != cudaSuccess) return e;
        const uint64_t n_max = max_items < uint64_t(num_sms) ? max_items : uint64_t(num_sms);  // one resident team per SM
        quant_general_team_kernel<<<unsigned(n_max ? n_max : 1), kGTeamWarps * 32, tsmem, st>>>(g, team_limit);
        e = cudaGetLastError();
        if (e != cudaSuccess) return e;
        if (!count_on_device) return cudaSuccess;
    }
    const uint64_t want = (max_items + kGWarps - 1) / kGWarps;
    const uint64_t cap = uint64_t(num_sms) * 3;  // 3 CTAs of ~70 KB shared memory per SM
    const unsigned grid = unsigned(want < cap ? (want ? want : 1) : cap);
    quant_general_kernel<<<grid, kGWarps * 32, smem, st>>>(g, team_limit);
    return cudaGetLastError();
}
