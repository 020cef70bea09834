// quant_covar.cuh — K3C / K4C: the quantitative regression WITH COVARIATES off the general kernel (included by
// kernels_quant.cu, inside namespace stoat).
//
// With covariates every snarl used to take K3G (one warp per snarl, ~170k warp instructions per snarl: 640 ms per 1M
// snarls at 5,000 samples).  The covariates cannot ride on the integer contraction of K3M as the phenotype does: the
// reference reads covariate row r for the r-th row of the FILTERED table (stats_test.cpp:450-452), so which covariate
// row meets which sample depends on the snarl as soon as one sample sits on no path.  What that indexing does allow:
//   * 1'C and C'C of X = [1 | dosages | C] are sums over the first n_used covariate rows — functions of n_used alone,
//     read from prefix tables made once per covariate matrix (abi.cu, stoat_gpu_set_covariates);
//   * what is left per snarl — n_used, 1'd, d'd, d'C, and X'y, y'y — takes ONE sweep over the samples.
// K3C does that sweep with a warp per snarl: the rows are loaded lane <-> word (coalesced), the arithmetic runs lane <->
// sample on 32 consecutive samples at a time (count fields handed over by shuffles, the rank of a sample among the used
// ones from a ballot), so that phenotype and covariate columns are read contiguously — a first version that kept a lane
// on its own 16 samples was bound by the L1's one-line-per-cycle rate (32 lines per load instruction: 75 ms per 1M snarls
// with 3 covariates, growing with their number).  K4C (a thread per snarl, as K4a) applies the filters
// (snarl_analyzer.cpp:542-579), completes X'X from the prefix tables, inverts it by the reference's rule (pivoted LDLT,
// ols.cuh) and takes rss from the moment expansion with the centred phenotype; the Student-t tails go to K4b.
// Snarls the formulas do not cover are pushed to the general list and K3G runs on those only: more than 4 paths or more
// than 2 checked edges on a path, an empty path (missing edge, no carrier), identical columns (the reference merges them,
// :620-674), a filter comparison within 1e-9 of its threshold with thirds in the table (K3G replays the reference's
// order of additions), a rank-deficient X'X (pseudo-inverse branch, not shift invariant), a near-perfect fit
// (rss <= 1e-4 tss: the expansion has lost its digits), more than kCovFeat features.
#pragma once

constexpr int kCovMaxD = 3;   // regression dosage columns (paths - 1)
constexpr int kCovMaxQ = 5;   // covariates
constexpr int kCovFeat = 8;   // features 1 + D + Q the solver takes
// rss from the moment expansion is trusted above this fraction of tss; below it the general kernel takes explicit
// residuals.  (1e-6, K4a's guard, let one pair in 29,040 randomised runs through at 1.3e-9: 5 samples, 1 - R^2 = 2e-6.)
constexpr double kCovRssGuard = 1e-4;
constexpr uint32_t kCovRowAll = 0xFFFFFFFEu;  // a path without checked edges: every haplotype

// per snarl, K3C -> K4C.  v[]: s1d[3] | sdd[6] (lower triangle) | sdy[3] | sy | syy | scy[Q] | sdc[3][Q]
struct CovMoments {
    double v[14 + kCovMaxQ + kCovMaxD * kCovMaxQ];
    uint32_t n_used;
    uint8_t n_dos;      // D = paths - 1
    uint8_t state;      // 0: nothing to solve here (on the general list), 1: moments valid
    uint8_t nondyadic;  // a sample on 3, 5, 6 or 7 path-haplotypes
    uint8_t pad;
};
constexpr int kCovOffSdd = 3, kCovOffSdy = 9, kCovOffSy = 12, kCovOffSyy = 13, kCovOffScy = 14;

struct CovParams {
    const uint8_t* rows;
    uint32_t pitch, n_hap, n_samples, n_covar;
    const double* yc;      // centred phenotype, sample order
    const double* covar_t;  // covariates by column: [n_covar][cov_rows]
    uint32_t cov_rows;
    const double* prefix;  // [n_samples + 1][n_covar + n_covar (n_covar + 1) / 2]: sums over the first r covariate rows
    BatchDev b;
    stoat_gpu_thresholds thr;
    QuantOut out;
    CovMoments* mom;
    uint32_t* g_list;
    uint32_t* g_count;
};

__device__ __forceinline__ void cov_push_general(const CovParams& c, uint64_t s) {
    const uint32_t slot = atomicAdd(c.g_count, 1u);
    c.g_list[slot] = uint32_t(s);
}

// count / row sum for counts 0..2 and row sums 0..8 (quantitative_table.cpp:93): k * rn(1 / rs) is rn(k / rs) for k <= 2
__device__ __forceinline__ void cov_fill_lut(double* lut, uint32_t tid, uint32_t n_threads) {
    for (uint32_t i = tid; i < 27; i += n_threads) {
        const uint32_t rs = i / 3, k = i % 3;
        lut[i] = (rs == 0 || k == 0) ? 0.0 : __ddiv_rn(double(k), double(rs));
    }
}

constexpr int kCovWarps = 8;

// The sweep of one snarl of NP paths (D = NP - 1 regression columns) by one warp: lane <-> words lane, lane + 32, ...
// Everything is sized by NP and Q at compile time, so a two-path snarl carries 11 accumulators, not 34.
template <int NP, int Q>
__device__ __forceinline__ void cov_sweep(const CovParams& c, const double* lut, uint64_t s, uint64_t p0,
                                          const uint32_t (&r0)[kCovMaxD + 1], const uint32_t (&r1)[kCovMaxD + 1],
                                          uint32_t lane) {
    constexpr int D = NP - 1;
    constexpr int kTri = D * (D + 1) / 2;
    const uint32_t n_words = c.pitch / 4;
    const uint32_t M = 0x55555555u;
    double s1d[D], sdd[kTri], sdy[D], sdc[D][Q], scy[Q], sy = 0.0, syy = 0.0;
#pragma unroll
    for (int a = 0; a < D; ++a) {
        s1d[a] = sdy[a] = 0.0;
#pragma unroll
        for (int j = 0; j < Q; ++j) sdc[a][j] = 0.0;
    }
#pragma unroll
    for (int e = 0; e < kTri; ++e) sdd[e] = 0.0;
#pragma unroll
    for (int j = 0; j < Q; ++j) scy[j] = 0.0;
    uint32_t al[NP];  // haplotypes on path j
#pragma unroll
    for (int j = 0; j < NP; ++j) al[j] = 0;
    uint32_t same = 0x3Fu;   // pair (a, b) of columns still identical: bit a (a - 1) / 2 + b, b < a
    uint32_t rs_seen = 0;    // bit rs set for every row sum met
    uint32_t rank_run = 0;   // used samples in the words of the earlier iterations
    const uint32_t src_half = lane >> 4, sh = 2 * (lane & 15), lt_mask = (1u << lane) - 1u;
    for (uint32_t w0 = 0; w0 < n_words; w0 += 32) {
        // a superblock: 32 words = 512 samples.  The words are loaded lane <-> word (one 128-byte request per row) ...
        const uint32_t w = w0 + lane;
        const bool in = w < n_words;
        uint32_t f[NP];  // per sample a 2-bit field: haplotypes of the sample on path j (0..2)
#pragma unroll
        for (int j = 0; j < NP; ++j) {
            uint32_t pw = 0u;
            if (in) {
                if (r0[j] == kCovRowAll) {
                    pw = valid_word(w, c.n_hap);
                } else {
                    const uint32_t va = __ldg(reinterpret_cast<const uint32_t*>(c.rows + size_t(r0[j]) * c.pitch) + w);
                    const uint32_t vb = __ldg(reinterpret_cast<const uint32_t*>(c.rows + size_t(r1[j]) * c.pitch) + w);
                    pw = va & vb;
                }
            }
            al[j] += __popc(pw);
            f[j] = pw - ((pw >> 1) & M);
        }
        if (NP >= 3) {  // identical columns (:620-674 merges them): the per-sample counts compared word-wise
            int bit = 0;
#pragma unroll
            for (int a = 1; a < NP; ++a)
#pragma unroll
                for (int b2 = 0; b2 < a; ++b2, ++bit)
                    if (f[a] != f[b2]) same &= ~(1u << bit);
        }
        // ... and worked lane <-> sample, 32 consecutive samples at a time (their count fields come by shuffle from the two
        // lanes that hold the block's words), so that the phenotype and the covariate columns are read contiguously
        const uint32_t n_blk = min(16u, (n_words - w0 + 1) / 2);
#pragma unroll 4
        for (uint32_t mm = 0; mm < n_blk; ++mm) {
            uint32_t k[NP], rs = 0;
#pragma unroll
            for (int j = 0; j < NP; ++j) {
                k[j] = (__shfl_sync(0xffffffffu, f[j], 2 * mm + src_half) >> sh) & 3u;
                rs += k[j];
            }
            const bool us = rs != 0;  // rs == 0: not a row of the table (quantitative_table.cpp:84-88)
            const uint32_t ballot = __ballot_sync(0xffffffffu, us);
            const uint32_t rank = rank_run + __popc(ballot & lt_mask);  // position in the FILTERED table
            rank_run += __popc(ballot);
            double y = 0.0, cv[Q];
#pragma unroll
            for (int j = 0; j < Q; ++j) cv[j] = 0.0;
            if (us) {
                y = __ldg(c.yc + size_t(w0) * 16 + mm * 32 + lane);
#pragma unroll
                for (int j = 0; j < Q; ++j) cv[j] = __ldg(c.covar_t + size_t(j) * c.cov_rows + rank);  // (stats_test.cpp:450-452)
            }
            rs_seen |= 1u << rs;
            double d[D];
#pragma unroll
            for (int a = 0; a < D; ++a) d[a] = lut[rs * 3 + k[a]];
            sy += y;
            syy = fma(y, y, syy);
#pragma unroll
            for (int j = 0; j < Q; ++j) scy[j] = fma(cv[j], y, scy[j]);
            int e = 0;
#pragma unroll
            for (int a = 0; a < D; ++a) {
                s1d[a] += d[a];
                sdy[a] = fma(d[a], y, sdy[a]);
#pragma unroll
                for (int j = 0; j < Q; ++j) sdc[a][j] = fma(d[a], cv[j], sdc[a][j]);
#pragma unroll
                for (int b2 = 0; b2 <= a; ++b2) sdd[e + b2] = fma(d[a], d[b2], sdd[e + b2]);
                e += a + 1;
            }
        }
    }
    // ---- reductions, hand-over -------------------------------------------------------------------------------------
    const uint32_t n_used = rank_run;
    bool empty_path = false, identical = false;
#pragma unroll
    for (int j = 0; j < NP; ++j) {
        al[j] = __reduce_add_sync(0xffffffffu, al[j]);
        if (lane == uint32_t(j)) c.out.allele_paths[p0 + j] = al[j];
        empty_path |= al[j] == 0;
    }
    if (NP >= 3) {
        same = __reduce_and_sync(0xffffffffu, same);
        identical = (same & ((1u << (NP * (NP - 1) / 2)) - 1u)) != 0u;
    }
    rs_seen = __reduce_or_sync(0xffffffffu, rs_seen);
    CovMoments* m = c.mom + s;
    if (empty_path || identical) {  // a column the reference drops or merges: the general kernel does that
        if (lane == 0) {
            m->state = 0;
            cov_push_general(c, s);
        }
        return;
    }
    double mine = 0.0;  // lane e keeps entry e of CovMoments::v
    auto put = [&](int e, double v) {
        const double tot = warp_sum(v);
        if (lane == uint32_t(e)) mine = tot;
    };
#pragma unroll
    for (int a = 0; a < D; ++a) put(a, s1d[a]);
#pragma unroll
    for (int e = 0; e < kTri; ++e) put(kCovOffSdd + e, sdd[e]);
#pragma unroll
    for (int a = 0; a < D; ++a) put(kCovOffSdy + a, sdy[a]);
    put(kCovOffSy, sy);
    put(kCovOffSyy, syy);
#pragma unroll
    for (int j = 0; j < Q; ++j) put(kCovOffScy + j, scy[j]);
    if (lane < uint32_t(kCovOffScy + Q)) m->v[lane] = mine;
    mine = 0.0;  // d'C: a second round of lanes
#pragma unroll
    for (int a = 0; a < D; ++a)
#pragma unroll
        for (int j = 0; j < Q; ++j) put(a * Q + j, sdc[a][j]);
    if (lane < uint32_t(D * Q)) m->v[kCovOffScy + Q + lane] = mine;
    if (lane == 0) {
        m->n_used = n_used;
        m->n_dos = uint8_t(D);
        m->state = 1;
        m->nondyadic = (rs_seen & 0xE8u) ? 1 : 0;  // row sums 3, 5, 6, 7
    }
}

// One kernel per path count (its registers are sized by NP and Q): the instance NP takes the snarls of NP paths; the
// two-path instance also turns away the shapes no instance takes.
template <int NP, int Q>
__global__ void __launch_bounds__(kCovWarps * 32, NP == 2 ? 3 : 2) covar_moments_kernel(const CovParams c) {
    __shared__ double lut[27];
    cov_fill_lut(lut, threadIdx.x, blockDim.x);
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (uint64_t s = uint64_t(blockIdx.x) * kCovWarps + warp; s < c.b.n_snarls; s += uint64_t(gridDim.x) * kCovWarps) {
        const uint64_t p0 = c.b.snarl_path_off[s];
        const uint32_t np = uint32_t(c.b.snarl_path_off[s + 1] - p0);
        // ---- the paths' rows: lane j reads path j (<= 2 checked edges, none missing), then broadcast -------------------
        bool shape_ok = np >= 2 && np <= uint32_t(kCovMaxD + 1) && 1 + (np - 1) + Q <= uint32_t(kCovFeat);
        if (np != uint32_t(NP) && (shape_ok || NP != 2)) continue;  // another instance's snarl
        uint32_t my_r0 = kCovRowAll, my_r1 = kCovRowAll;
        if (shape_ok && lane < np) {
            const uint64_t a = c.b.path_edge_off[p0 + lane], z = c.b.path_edge_off[p0 + lane + 1];
            if (z - a > 2) shape_ok = false;
            else if (z - a >= 1) {
                my_r0 = c.b.edge_row[a];
                my_r1 = (z - a == 2) ? c.b.edge_row[a + 1] : my_r0;
                if (my_r0 == kMissingQ || my_r1 == kMissingQ) shape_ok = false;
            }
        }
        shape_ok = __all_sync(0xffffffffu, shape_ok);
        if (!shape_ok) {  // (the two-path instance for a path count nobody takes, any instance for its own snarls)
            if (lane == 0) {
                c.mom[s].state = 0;
                cov_push_general(c, s);
            }
            continue;
        }
        uint32_t r0[kCovMaxD + 1], r1[kCovMaxD + 1];
#pragma unroll
        for (int j = 0; j <= kCovMaxD; ++j) {
            r0[j] = __shfl_sync(0xffffffffu, my_r0, j);
            r1[j] = __shfl_sync(0xffffffffu, my_r1, j);
        }
        cov_sweep<NP, Q>(c, lut, s, p0, r0, r1, lane);
    }
}

// ---- K4C: filters, X'X completed from the prefix tables, solve, statistics (a thread per snarl) ------------------------
__global__ void __launch_bounds__(128) covar_solve_kernel(const CovParams c, const SolveArgs a) {
    const uint64_t s = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (s >= c.b.n_snarls) return;
    const CovMoments& m = c.mom[s];
    if (m.state != 1) return;  // on the general list
    const uint32_t Q = c.n_covar, D = m.n_dos, nk = D + 1, n_used = m.n_used;
    // filtration_quantitative_table (snarl_analyzer.cpp:542-579) as quant_general.cuh applies it
    if (n_used == 0 || uint64_t(n_used) < c.thr.min_individuals) {
        a.out.filtered[s] = 1;
        a.out.n_used[s] = n_used;
        return;  // p / beta / se / r2 = NaN, code NA: preset by the caller
    }
    double colsum[kCovMaxD + 1], last = double(n_used);
    for (uint32_t j = 0; j < D; ++j) {
        colsum[j] = m.v[j];
        last -= m.v[j];  // the dosages of a row add up to 1: exact for dyadic dosages
    }
    colsum[D] = last;
    const double tot = double(n_used), mh = double(c.thr.min_haplotypes), eps = 1e-9;
    bool borderline = fabs(tot - mh) <= eps * fmax(tot, 1.0);
    int above = 0;
    for (uint32_t j = 0; j < nk; ++j) {
        const double freq = __ddiv_rn(colsum[j], tot);
        const double maf = fmin(freq, __dsub_rn(1.0, freq));
        if (fabs(maf - c.thr.maf_threshold) <= eps) borderline = true;
        if (maf > c.thr.maf_threshold) ++above;
    }
    if (m.nondyadic && borderline) {  // the outcome hangs on the reference's order of additions: K3G replays it
        solve_hand_on(a, s);
        return;
    }
    if (tot < mh || above < 2) {
        a.out.filtered[s] = 1;
        a.out.n_used[s] = n_used;
        return;
    }
    a.out.filtered[s] = 0;
    a.out.n_used[s] = n_used;
    // X = [1 | d_0 .. d_(D-1) | c_0 .. c_(Q-1)] (the last dosage column is dropped, :676-684)
    const int p = int(1 + D + Q);
    const double* pre = c.prefix + size_t(n_used) * (Q + Q * (Q + 1) / 2);
    SmallMat<kCovFeat> X, inv;
    double xty[kCovFeat];
    X(0, 0) = double(n_used);
    xty[0] = m.v[kCovOffSy];
    for (uint32_t i = 0; i < D; ++i) {
        X(1 + i, 0) = X(0, 1 + i) = m.v[i];
        xty[1 + i] = m.v[kCovOffSdy + i];
        for (uint32_t j = 0; j <= i; ++j) X(1 + i, 1 + j) = X(1 + j, 1 + i) = m.v[kCovOffSdd + i * (i + 1) / 2 + j];
    }
    uint32_t e = Q;
    for (uint32_t j = 0; j < Q; ++j) {
        X(1 + D + j, 0) = X(0, 1 + D + j) = pre[j];
        xty[1 + D + j] = m.v[kCovOffScy + j];
        for (uint32_t i = 0; i < D; ++i) X(1 + D + j, 1 + i) = X(1 + i, 1 + D + j) = m.v[kCovOffScy + Q + i * Q + j];
        for (uint32_t k = j; k < Q; ++k, ++e) X(1 + D + j, 1 + D + k) = X(1 + D + k, 1 + D + j) = pre[e];
    }
    if (xtx_inverse<kCovFeat>(X, p, inv)) {  // rank deficient: the pseudo-inverse branch is not shift invariant
        solve_hand_on(a, s);
        return;
    }
    double beta[kCovFeat];
    for (int i = 0; i < p; ++i) {
        double b = 0.0;
        for (int j = 0; j < p; ++j) b += inv(i, j) * xty[j];
        beta[i] = b;
    }
    double bxy = 0.0, bxxb = 0.0;
    for (int i = 0; i < p; ++i) {
        bxy += beta[i] * xty[i];
        double r = 0.0;
        for (int j = 0; j < p; ++j) r += X(i, j) * beta[j];
        bxxb += beta[i] * r;
    }
    const double syy = m.v[kCovOffSyy], sy = m.v[kCovOffSy];
    const double rss = syy - 2.0 * bxy + bxxb;
    const double tss = syy - sy * sy / double(n_used);
    if (!(rss > kCovRssGuard * tss)) {  // near-perfect fit: explicit residuals on the general path
        solve_hand_on(a, s);
        return;
    }
    solve_store(a, s, p, n_used, beta[1], inv(1, 1), rss, tss);
}

size_t quant_covar_moments_bytes(uint64_t n_snarls) { return size_t(n_snarls) * sizeof(CovMoments); }
bool quant_covar_fast(uint32_t n_covar) {
    const char* v = getenv("STOAT_GPU_COVAR_FAST");  // 0: every snarl on the general kernel (tests compare the two)
    const bool off = v && v[0] == '0';
    return !off && n_covar >= 1 && n_covar <= uint32_t(kCovMaxQ);
}

template <int Q>
inline cudaError_t launch_covar_moments_q(const CovParams& c, unsigned grid, cudaStream_t st) {
    covar_moments_kernel<2, Q><<<grid, kCovWarps * 32, 0, st>>>(c);
    if constexpr (1 + 2 + Q <= kCovFeat) covar_moments_kernel<3, Q><<<grid, kCovWarps * 32, 0, st>>>(c);
    if constexpr (1 + 3 + Q <= kCovFeat) covar_moments_kernel<4, Q><<<grid, kCovWarps * 32, 0, st>>>(c);
    return cudaGetLastError();
}

inline cudaError_t launch_covar_moments(const CovParams& c, int num_sms, cudaStream_t st) {
    const uint64_t want = (c.b.n_snarls + kCovWarps - 1) / kCovWarps;
    const uint64_t cap = uint64_t(num_sms) * 12;
    const unsigned grid = unsigned(want < cap ? (want ? want : 1) : cap);
    switch (c.n_covar) {
        case 1: return launch_covar_moments_q<1>(c, grid, st);
        case 2: return launch_covar_moments_q<2>(c, grid, st);
        case 3: return launch_covar_moments_q<3>(c, grid, st);
        case 4: return launch_covar_moments_q<4>(c, grid, st);
        case 5: return launch_covar_moments_q<5>(c, grid, st);
        default: return cudaErrorInvalidValue;
    }
}

// =================================================================================================
// E3C: the eQTL regressions of gene windows, and of any scan WITH COVARIATES, off the general kernel (EQTLSnarlAnalyzer::analyze_and_write_snarl,
// src/snarl_analyzer.cpp:493-539, one OLS per gene of the snarl's window on the same table).
// A warp takes (snarl, share of its genes): one sweep over the samples for the design's own moments (as K3C) and, on lane
// 0, the filters and (X'X)^-1 — once per snarl, as the reference should have done (it re-inverts per gene, :510-520) —
// then one sweep per batch of GB genes for X'y_g and y_g'y_g (the expression rows are read contiguously, lane <-> sample)
// and lane g finishes gene g.  The phenotype is shifted by the mean of the gene's first 32 samples before it is squared
// (OLS with an intercept is shift invariant; the raw moments would cancel).  Tried and dropped: the expression values staged
// once per CTA in shared memory for its 8 snarls (cp.async double buffering with CTA barriers: 57 ms; a 4-stage TMA ring with
// full / empty mbarriers and no CTA barrier: 50 ms, against 44 ms for this form on 2,000 x 2,000 x 5,000) — the expression
// loads are not what a pair waits for; the per-block chain shuffle -> ballot -> rank -> covariate loads is.  Shapes K3C turns away, a rank-deficient
// design and any pair whose rss is below 1e-4 tss flag the whole snarl for the general kernel (flag[entry] = 1).
// =================================================================================================
struct CovEqtlParams {
    CovParams c;                 // rows, covariates, thresholds, out (per-pair arrays), b; yc / prefix / mom unused
    const double* prefix;        // as CovParams::prefix
    const double* expression;    // [n_genes][n_samples] raw
    uint32_t n_genes;
    const uint64_t* snarl_gene_off;  // nullptr: dense (the items of entry e are genes 0 .. n_genes - 1)
    const uint32_t* gene_idx;
    const uint32_t* list;        // nullptr: snarls s_begin + entry
    uint64_t s_begin, n_entries, item_base;
    uint32_t chunks;             // warps per entry, each a contiguous share of its items
    uint8_t* flag;               // [n_entries] 1: the general kernel takes the entry
};

template <int NP, int Q, typename Body>
__device__ __forceinline__ uint32_t cov_for_blocks(const CovParams& c, const uint32_t (&r0)[kCovMaxD + 1],
                                                   const uint32_t (&r1)[kCovMaxD + 1], uint32_t lane, uint32_t (&al)[NP],
                                                   uint32_t& same, Body&& body) {
    const uint32_t n_words = c.pitch / 4, M = 0x55555555u;
    const uint32_t src_half = lane >> 4, sh = 2 * (lane & 15), lt_mask = (1u << lane) - 1u;
    uint32_t rank_run = 0;
    for (uint32_t w0 = 0; w0 < n_words; w0 += 32) {
        const uint32_t w = w0 + lane;
        const bool in = w < n_words;
        uint32_t f[NP];
#pragma unroll
        for (int j = 0; j < NP; ++j) {
            uint32_t pw = 0u;
            if (in) {
                if (r0[j] == kCovRowAll) {
                    pw = valid_word(w, c.n_hap);
                } else {
                    const uint32_t va = __ldg(reinterpret_cast<const uint32_t*>(c.rows + size_t(r0[j]) * c.pitch) + w);
                    const uint32_t vb = __ldg(reinterpret_cast<const uint32_t*>(c.rows + size_t(r1[j]) * c.pitch) + w);
                    pw = va & vb;
                }
            }
            al[j] += __popc(pw);
            f[j] = pw - ((pw >> 1) & M);
        }
        if (NP >= 3) {
            int bit = 0;
#pragma unroll
            for (int a = 1; a < NP; ++a)
#pragma unroll
                for (int b2 = 0; b2 < a; ++b2, ++bit)
                    if (f[a] != f[b2]) same &= ~(1u << bit);
        }
        const uint32_t n_blk = min(16u, (n_words - w0 + 1) / 2);
#pragma unroll 2
        for (uint32_t mm = 0; mm < n_blk; ++mm) {
            uint32_t k[NP], rs = 0;
#pragma unroll
            for (int j = 0; j < NP; ++j) {
                k[j] = (__shfl_sync(0xffffffffu, f[j], 2 * mm + src_half) >> sh) & 3u;
                rs += k[j];
            }
            const bool us = rs != 0;
            const uint32_t ballot = __ballot_sync(0xffffffffu, us);
            const uint32_t rank = rank_run + __popc(ballot & lt_mask);
            rank_run += __popc(ballot);
            body(w0 * 16 + mm * 32 + lane, us, rs, k, rank);
        }
    }
    return rank_run;  // n_used
}

// The design of one snarl (one sweep, then lane 0: filters, X'X, inverse into wsm = [X'X (P x P) | inverse (P x P)]).
// -> 0: regress, 1: filtered, 2: the general kernel takes the snarl; allele_paths are written here.
template <int NP, int Q>
__device__ __forceinline__ int cov_eqtl_design(const CovEqtlParams& q, const double* lut, double* wsm, uint64_t s, uint64_t p0,
                                               const uint32_t (&r0)[kCovMaxD + 1], const uint32_t (&r1)[kCovMaxD + 1],
                                               uint32_t lane, uint32_t& n_used_out) {
    constexpr int D = NP - 1, kTri = D * (D + 1) / 2, P = 1 + D + Q;
    const CovParams& c = q.c;
    // ---- the design: one sweep --------------------------------------------------------------------------------------
    constexpr int QA = Q > 0 ? Q : 1;  // (Q = 0: the same kernels without covariates)
    double s1d[D], sdd[kTri], sdc[D][QA];
#pragma unroll
    for (int a = 0; a < D; ++a) {
        s1d[a] = 0.0;
#pragma unroll
        for (int j = 0; j < Q; ++j) sdc[a][j] = 0.0;
    }
#pragma unroll
    for (int e = 0; e < kTri; ++e) sdd[e] = 0.0;
    uint32_t al[NP];
#pragma unroll
    for (int j = 0; j < NP; ++j) al[j] = 0;
    uint32_t same = 0x3Fu, rs_seen = 0;
    const uint32_t n_used = cov_for_blocks<NP, Q>(c, r0, r1, lane, al, same,
        [&](uint32_t, bool us, uint32_t rs, const uint32_t (&k)[NP], uint32_t rank) {
            rs_seen |= 1u << rs;
            double cv[QA];
#pragma unroll
            for (int j = 0; j < Q; ++j) cv[j] = us ? __ldg(c.covar_t + size_t(j) * c.cov_rows + rank) : 0.0;
            int e = 0;
#pragma unroll
            for (int a = 0; a < D; ++a) {
                const double d = lut[rs * 3 + k[a]];
                s1d[a] += d;
#pragma unroll
                for (int j = 0; j < Q; ++j) sdc[a][j] = fma(d, cv[j], sdc[a][j]);
#pragma unroll
                for (int b2 = 0; b2 <= a; ++b2) sdd[e + b2] = fma(d, lut[rs * 3 + k[b2]], sdd[e + b2]);
                e += a + 1;
            }
        });
    bool empty_path = false, identical = false;
#pragma unroll
    for (int j = 0; j < NP; ++j) {
        al[j] = __reduce_add_sync(0xffffffffu, al[j]);
        if (lane == uint32_t(j)) c.out.allele_paths[p0 + j] = al[j];
        empty_path |= al[j] == 0;
    }
    if (NP >= 3) {
        same = __reduce_and_sync(0xffffffffu, same);
        identical = (same & ((1u << (NP * (NP - 1) / 2)) - 1u)) != 0u;
    }
    rs_seen = __reduce_or_sync(0xffffffffu, rs_seen);
    n_used_out = n_used;
    if (empty_path || identical) return 2;
#pragma unroll
    for (int a = 0; a < D; ++a) {
        s1d[a] = warp_sum(s1d[a]);
#pragma unroll
        for (int j = 0; j < Q; ++j) sdc[a][j] = warp_sum(sdc[a][j]);
    }
#pragma unroll
    for (int e = 0; e < kTri; ++e) sdd[e] = warp_sum(sdd[e]);
    // ---- lane 0: filters (snarl_analyzer.cpp:542-579), X'X, its inverse; wsm = [X'X (P x P) | inverse (P x P)] ------------
    int status = 0;  // 0: regress, 1: filtered, 2: general kernel
    if (lane == 0) {
        if (n_used == 0 || uint64_t(n_used) < c.thr.min_individuals) {
            status = 1;
        } else {
            double colsum[NP], last = double(n_used);
#pragma unroll
            for (int j = 0; j < D; ++j) {
                colsum[j] = s1d[j];
                last -= s1d[j];
            }
            colsum[D] = last;
            const double tot = double(n_used), mh = double(c.thr.min_haplotypes), eps = 1e-9;
            bool borderline = fabs(tot - mh) <= eps * fmax(tot, 1.0);
            int above = 0;
#pragma unroll
            for (int j = 0; j < NP; ++j) {
                const double freq = __ddiv_rn(colsum[j], tot);
                const double maf = fmin(freq, __dsub_rn(1.0, freq));
                if (fabs(maf - c.thr.maf_threshold) <= eps) borderline = true;
                if (maf > c.thr.maf_threshold) ++above;
            }
            if ((rs_seen & 0xE8u) && borderline) status = 2;
            else if (tot < mh || above < 2) status = 1;
        }
        if (status == 0) {
            const double* pre = q.prefix + size_t(n_used) * (Q + Q * (Q + 1) / 2);
            SmallMat<kCovFeat> X, inv;
            X(0, 0) = double(n_used);
            for (int i = 0; i < D; ++i) {
                X(1 + i, 0) = X(0, 1 + i) = s1d[i];
                for (int j = 0; j <= i; ++j) X(1 + i, 1 + j) = X(1 + j, 1 + i) = sdd[i * (i + 1) / 2 + j];
            }
            int e = Q;
            for (int j = 0; j < Q; ++j) {
                X(1 + D + j, 0) = X(0, 1 + D + j) = pre[j];
                for (int i = 0; i < D; ++i) X(1 + D + j, 1 + i) = X(1 + i, 1 + D + j) = sdc[i][j];
                for (int k2 = j; k2 < Q; ++k2, ++e) X(1 + D + j, 1 + D + k2) = X(1 + D + k2, 1 + D + j) = pre[e];
            }
            if (xtx_inverse<kCovFeat>(X, P, inv)) {
                status = 2;  // rank deficient: the pseudo-inverse branch is not shift invariant
            } else {
                for (int i = 0; i < P; ++i)
                    for (int j = 0; j < P; ++j) {
                        wsm[i * P + j] = X(i, j);
                        wsm[P * P + i * P + j] = inv(i, j);
                    }
            }
        }
    }
    status = __shfl_sync(0xffffffffu, status, 0);
    __syncwarp();
    return status;
}

// One gene of a snarl from its X'y (lane-private): beta, rss by the moment expansion, statistics, stored at item oi.
// -> false: the expansion has lost its digits (the general kernel takes explicit residuals).
template <int P>
__device__ __forceinline__ bool cov_eqtl_finish(const CovParams& c, const double* wsm, const double (&xty)[P], double my_syy,
                                                uint32_t n_used, uint64_t oi) {
    double beta[P];
#pragma unroll
    for (int i = 0; i < P; ++i) {
        double b = 0.0;
#pragma unroll
        for (int j = 0; j < P; ++j) b += wsm[P * P + i * P + j] * xty[j];
        beta[i] = b;
    }
    double bxy = 0.0, bxxb = 0.0;
#pragma unroll
    for (int i = 0; i < P; ++i) {
        bxy += beta[i] * xty[i];
        double r = 0.0;
#pragma unroll
        for (int j = 0; j < P; ++j) r += wsm[i * P + j] * beta[j];
        bxxb += beta[i] * r;
    }
    const double rss = my_syy - 2.0 * bxy + bxxb;
    const double tss = my_syy - xty[0] * xty[0] / double(n_used);
    if (!(rss > kCovRssGuard * tss)) return false;  // near-perfect fit: explicit residuals, on the general path
    const OlsOut o = ols_statistics(beta[1], wsm[P * P + P + 1], rss, tss, n_used, P);
    c.out.p[oi] = o.p;
    c.out.r2[oi] = o.r2;
    c.out.beta[oi] = o.beta;
    c.out.se[oi] = o.se;
    c.out.code[oi] = kPValue;
    return true;
}

template <int NP, int Q, int GB>
__device__ __forceinline__ void cov_eqtl_snarl(const CovEqtlParams& q, const double* lut, double* wsm, uint64_t entry,
                                               uint64_t s, uint32_t part, uint64_t p0, const uint32_t (&r0)[kCovMaxD + 1],
                                               const uint32_t (&r1)[kCovMaxD + 1], uint32_t lane) {
    constexpr int D = NP - 1, kTri = D * (D + 1) / 2, P = 1 + D + Q, QA = Q > 0 ? Q : 1;
    const CovParams& c = q.c;
    const double nan = __longlong_as_double(0x7ff8000000000000LL);
    // ---- the items (genes) of this warp ------------------------------------------------------------------------------
    const uint64_t dense_row = q.list ? entry : s;  // as quant_general.cuh numbers the items of a dense scan
    uint64_t k0 = q.snarl_gene_off ? q.snarl_gene_off[s] : dense_row * uint64_t(q.n_genes);
    uint64_t k1 = q.snarl_gene_off ? q.snarl_gene_off[s + 1] : (dense_row + 1) * uint64_t(q.n_genes);
    const uint64_t gene_base = dense_row * uint64_t(q.n_genes);
    {
        const uint64_t share = (k1 - k0 + q.chunks - 1) / q.chunks;
        const uint64_t lo = k0 + uint64_t(part) * share;
        k1 = lo + share < k1 ? lo + share : k1;
        k0 = lo < k1 ? lo : k1;
    }
    uint32_t n_used = 0;
    int status = cov_eqtl_design<NP, Q>(q, lut, wsm, s, p0, r0, r1, lane, n_used);
    if (status == 2) {
        if (lane == 0) q.flag[entry] = 1;
        return;
    }
    if (lane == 0) {
        c.out.filtered[s] = status == 1 ? 1 : 0;
        c.out.n_used[s] = n_used;
    }
    if (status == 1) {
        for (uint64_t k = k0 + lane; k < k1; k += 32) {
            const uint64_t o = k - q.item_base;
            c.out.p[o] = c.out.beta[o] = c.out.se[o] = c.out.r2[o] = nan;
            c.out.code[o] = kPNa;
        }
        return;
    }
    // ---- the genes, GB at a time --------------------------------------------------------------------------------------
    bool redo = false;
    for (uint64_t kb = k0; kb < k1; kb += GB) {
        const double* yrow[GB];
        double shift[GB];
#pragma unroll
        for (int g = 0; g < GB; ++g) {
            const uint64_t k = kb + g < k1 ? kb + g : k1 - 1;  // (a short last batch repeats its last gene)
            const uint64_t gene = q.gene_idx ? uint64_t(q.gene_idx[k]) : (k - gene_base);
            yrow[g] = q.expression + gene * uint64_t(c.n_samples);
            const double v = lane < c.n_samples ? __ldg(yrow[g] + lane) : 0.0;
            shift[g] = warp_sum(v) / double(c.n_samples < 32u ? c.n_samples : 32u);
        }
        double sy[GB], syy[GB], sdy[GB][D], scy[GB][QA];
#pragma unroll
        for (int g = 0; g < GB; ++g) {
            sy[g] = syy[g] = 0.0;
#pragma unroll
            for (int a = 0; a < D; ++a) sdy[g][a] = 0.0;
#pragma unroll
            for (int j = 0; j < Q; ++j) scy[g][j] = 0.0;
        }
        uint32_t al2[NP];
#pragma unroll
        for (int j = 0; j < NP; ++j) al2[j] = 0;
        uint32_t same2 = 0;
        cov_for_blocks<NP, Q>(c, r0, r1, lane, al2, same2,
            [&](uint32_t i, bool us, uint32_t rs, const uint32_t (&k)[NP], uint32_t rank) {
                double cv[QA], d[D];
#pragma unroll
                for (int j = 0; j < Q; ++j) cv[j] = us ? __ldg(c.covar_t + size_t(j) * c.cov_rows + rank) : 0.0;
#pragma unroll
                for (int a = 0; a < D; ++a) d[a] = lut[rs * 3 + k[a]];
#pragma unroll
                for (int g = 0; g < GB; ++g) {
                    const double y = us ? __ldg(yrow[g] + i) - shift[g] : 0.0;
                    sy[g] += y;
                    syy[g] = fma(y, y, syy[g]);
#pragma unroll
                    for (int a = 0; a < D; ++a) sdy[g][a] = fma(d[a], y, sdy[g][a]);
#pragma unroll
                    for (int j = 0; j < Q; ++j) scy[g][j] = fma(cv[j], y, scy[g][j]);
                }
            });
        // lane g finishes gene g: X'y = [sy, sdy, scy]
        double xty[P], my_syy = 0.0;
#pragma unroll
        for (int i = 0; i < P; ++i) xty[i] = 0.0;
#pragma unroll
        for (int g = 0; g < GB; ++g) {
            const double t_sy = warp_sum(sy[g]), t_syy = warp_sum(syy[g]);
            if (lane == uint32_t(g)) {
                xty[0] = t_sy;
                my_syy = t_syy;
            }
#pragma unroll
            for (int a = 0; a < D; ++a) {
                const double t = warp_sum(sdy[g][a]);
                if (lane == uint32_t(g)) xty[1 + a] = t;
            }
#pragma unroll
            for (int j = 0; j < Q; ++j) {
                const double t = warp_sum(scy[g][j]);
                if (lane == uint32_t(g)) xty[1 + D + j] = t;
            }
        }
        bool bad = false;
        if (lane < uint32_t(GB) && kb + lane < k1) bad = !cov_eqtl_finish<P>(c, wsm, xty, my_syy, n_used, kb + lane - q.item_base);
        if (__any_sync(0xffffffffu, bad)) {
            redo = true;
            break;
        }
    }
    if (redo && lane == 0) q.flag[entry] = 1;
}

constexpr int kCovEqtlWarps = 8;

template <int NP, int Q, int GB, int MINB>
__global__ void __launch_bounds__(kCovEqtlWarps * 32, MINB) eqtl_covar_kernel(const CovEqtlParams q) {
    __shared__ double lut[27];
    __shared__ double wsm_all[kCovEqtlWarps][2 * kCovFeat * kCovFeat];
    cov_fill_lut(lut, threadIdx.x, blockDim.x);
    __syncthreads();
    const CovParams& c = q.c;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint64_t n_work = q.n_entries * q.chunks;
    for (uint64_t i = uint64_t(blockIdx.x) * kCovEqtlWarps + warp; i < n_work; i += uint64_t(gridDim.x) * kCovEqtlWarps) {
        // consecutive warps (the warps of a CTA) take consecutive snarls and the SAME share of the genes: they walk the same
        // expression rows at about the same time, so all but the first find them in L1 (a pair re-reads its gene's whole
        // row: with every warp on its own genes the scan was bound by the L2 -> SM traffic, 2.8 TB/s)
        const uint32_t part = uint32_t(i / q.n_entries);
        const uint64_t entry = i - uint64_t(part) * q.n_entries;
        const uint64_t s = q.list ? uint64_t(q.list[entry]) : q.s_begin + entry;
        const uint64_t p0 = c.b.snarl_path_off[s];
        const uint32_t np = uint32_t(c.b.snarl_path_off[s + 1] - p0);
        bool shape_ok = np >= 2 && np <= uint32_t(kCovMaxD + 1) && 1 + (np - 1) + Q <= uint32_t(kCovFeat);
        if (np != uint32_t(NP) && (shape_ok || NP != 2)) continue;  // another instance's snarl
        uint32_t my_r0 = kCovRowAll, my_r1 = kCovRowAll;
        if (shape_ok && lane < np) {
            const uint64_t a = c.b.path_edge_off[p0 + lane], z = c.b.path_edge_off[p0 + lane + 1];
            if (z - a > 2) shape_ok = false;
            else if (z - a >= 1) {
                my_r0 = c.b.edge_row[a];
                my_r1 = (z - a == 2) ? c.b.edge_row[a + 1] : my_r0;
                if (my_r0 == kMissingQ || my_r1 == kMissingQ) shape_ok = false;
            }
        }
        shape_ok = __all_sync(0xffffffffu, shape_ok);
        if (!shape_ok) {
            if (lane == 0) q.flag[entry] = 1;
            continue;
        }
        uint32_t r0[kCovMaxD + 1], r1[kCovMaxD + 1];
#pragma unroll
        for (int j = 0; j <= kCovMaxD; ++j) {
            r0[j] = __shfl_sync(0xffffffffu, my_r0, j);
            r1[j] = __shfl_sync(0xffffffffu, my_r1, j);
        }
        cov_eqtl_snarl<NP, Q, GB>(q, lut, wsm_all[warp], entry, s, part, p0, r0, r1, lane);
        __syncwarp();
    }
}

template <int Q>
inline cudaError_t launch_eqtl_covar_q(const CovEqtlParams& q, unsigned grid, cudaStream_t st) {
    static const int variant = [] {
        const char* v = getenv("STOAT_GPU_E3C");
        return v ? atoi(v) : 0;
    }();
    if (variant == 1) eqtl_covar_kernel<2, Q, 4, 2><<<grid, kCovEqtlWarps * 32, 0, st>>>(q);
    else eqtl_covar_kernel<2, Q, 8, 1><<<grid, kCovEqtlWarps * 32, 0, st>>>(q);
    if constexpr (1 + 2 + Q <= kCovFeat) eqtl_covar_kernel<3, Q, 4, 1><<<grid, kCovEqtlWarps * 32, 0, st>>>(q);
    if constexpr (1 + 3 + Q <= kCovFeat) eqtl_covar_kernel<4, Q, 4, 1><<<grid, kCovEqtlWarps * 32, 0, st>>>(q);
    return cudaGetLastError();
}

inline cudaError_t launch_eqtl_covar(CovEqtlParams& q, uint64_t max_items_per_entry, int num_sms, cudaStream_t st) {
    // warps per entry: enough to fill the machine, never less than 64 genes (8 batches) per warp unless the entry is short
    const uint64_t target = uint64_t(num_sms) * kCovEqtlWarps * 8;  // ~8 work items per resident warp: an even last round
    const uint64_t want = (target + q.n_entries - 1) / q.n_entries;
    const uint64_t most = std::max<uint64_t>(1, max_items_per_entry / 64);
    q.chunks = uint32_t(std::max<uint64_t>(1, std::min(want, most)));
    const uint64_t n_work = q.n_entries * q.chunks;
    const uint64_t cap = uint64_t(num_sms) * 8;
    const uint64_t ctas = (n_work + kCovEqtlWarps - 1) / kCovEqtlWarps;
    const unsigned grid = unsigned(ctas < cap ? ctas : cap);
    switch (q.c.n_covar) {
        case 0: return launch_eqtl_covar_q<0>(q, grid, st);
        case 1: return launch_eqtl_covar_q<1>(q, grid, st);
        case 2: return launch_eqtl_covar_q<2>(q, grid, st);
        case 3: return launch_eqtl_covar_q<3>(q, grid, st);
        case 4: return launch_eqtl_covar_q<4>(q, grid, st);
        case 5: return launch_eqtl_covar_q<5>(q, grid, st);
        default: return cudaErrorInvalidValue;
    }
}

