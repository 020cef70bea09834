// quant_reg.cuh — K3R, the register-resident form of K3 for two-path snarls (included by kernels_quant.cu,
// namespace stoat).
//
// K3 (quant_lean.cuh) gives a snarl to a warp, so a lane meets different samples in every iteration and has to fetch
// their phenotype values from shared memory: one LDS.64 + bit tests + selects per sample and class, ~20 instructions per
// sample, which made the quantitative step issue bound at 0.14 of the HBM roofline.  K3R turns the mapping round: a CTA
// takes ONE snarl at a time and thread t always owns the same WPT 32-haplotype words (16 * WPT samples), so the centred
// phenotype of those samples is loaded ONCE into registers and stays there for the whole kernel.  Per sample and class
// the inner loop is then a predicate-writing LOP3 and a predicated DADD — 4 issue slots per sample, at the rate of the
// FP64 pipe (2 DADDs per sample).
//
//   * rows are staged exactly as in K1 / K3 (cp.async.bulk into a ring of shared-memory stages, one mbarrier per stage,
//     issued lane-distributed by warp 0), but the ring belongs to the CTA; a stage is refilled after the CTA-wide
//     barrier that also publishes the per-warp partial sums;
//   * per snarl every warp reduces its class counts (packed REDUX) and three FP64 sums (shuffles), a rotating warp adds
//     the warps' partials and either finishes the snarl (filters of snarl_analyzer.cpp:542-579, moments for K4) or, when
//     the haplotype range is split over gridDim.y slices / over GPUs (sample sharding), stores the slice's partial sums
//     for K3F (lean_finish_kernel), which runs after the slices are added up;
//   * any other shape (1, 3, 4 paths, > 4 checked edges) goes to the slow list and is done by K3's warp-per-snarl kernel.
#pragma once

struct LeanSums {
    uint32_t al0, al1, n_used, n_full0, n_half, n_third, n_tthird;
    double xty1, su_un, suu_un;  // sum d0 * y over the used samples; sum y, sum y^2 over the UNUSED samples
};

// SoA partial sums of a slice (and the buffer a sample-sharded run all-reduces): ints [7][S], doubles [3][S]
struct LeanPartials {
    uint32_t* i;
    double* d;
    uint64_t stride;  // S
};

// n_tthird partial >= this: some slice met a sample on three path-haplotypes and did not sum it (K3M) -> general path
constexpr uint32_t kLeanReplayMark = 0x01000000u;

struct K3RParams {
    K3Params k;
    uint32_t slice_words;  // words of a row one CTA (slice) covers; == n_words when gridDim.y == 1
    uint32_t n_ring;       // stages in the CTA's ring
    uint32_t* slow_list;
    uint32_t* slow_count;
    LeanPartials part;     // used when partial_only
    uint32_t partial_only;
    uint32_t zero4[4];  // zeros (see dosage_weight_at)
};

// Filters + moments of a two-path snarl from its class counts and sums (what quant_lean2 does after its passes).
// One thread.  Returns 0: filtered (outputs written), 1: moments ready in `mom`, 2: the snarl must be replayed by the
// general path (non-dyadic borderline filters).
// force: -1 decide here; 0 / 1: the decision of the total / allele-frequency filters was taken by an exact replay of the
// reference's summation order (lean_exact_filtered, quant_mma.cuh): 1 = filtered
__device__ int lean_finish(const K3Params& p, uint64_t s, uint32_t p0, const LeanSums& t, QuantMoments& mom,
                           int force = -1) {
    p.out.allele_paths[p0] = t.al0;  // quantitative_table.cpp:54-55
    p.out.allele_paths[p0 + 1] = t.al1;
    if (t.n_tthird >= kLeanReplayMark) return 2;
    const bool nondyadic = (t.n_third | t.n_tthird) != 0u;
    const uint32_t n_used = t.n_used;
    const uint32_t n_full1 = n_used - t.n_full0 - t.n_half - t.n_third - t.n_tthird;
    const double third = 1.0 / 3.0, tthird = 2.0 / 3.0;
    const double colsum0 =
        double(t.n_full0) + 0.5 * double(t.n_half) + third * double(t.n_third) + tthird * double(t.n_tthird);
    const double colsum1 = double(n_full1) + 0.5 * double(t.n_half) + tthird * double(t.n_third) + third * double(t.n_tthird);
    bool filtered = (n_used == 0) || (t.al0 == 0) || (t.al1 == 0) || (uint64_t(n_used) < p.thr.min_individuals);
    if (!filtered && force >= 0) {
        filtered = force != 0;
    } else if (!filtered) {
        const double total = double(n_used);  // every used row sums to 1
        if (nondyadic && uint64_t(n_used) == p.thr.min_haplotypes) return 2;  // borderline: exact replay
        if (total < double(p.thr.min_haplotypes)) filtered = true;
        const double f0 = __ddiv_rn(colsum0, total), f1 = __ddiv_rn(colsum1, total);
        const double maf0 = fmin(f0, __dsub_rn(1.0, f0)), maf1 = fmin(f1, __dsub_rn(1.0, f1));
        if (nondyadic && (fabs(maf0 - p.thr.maf_threshold) <= 1e-9 || fabs(maf1 - p.thr.maf_threshold) <= 1e-9))
            return 2;
        if (!(maf0 > p.thr.maf_threshold && maf1 > p.thr.maf_threshold)) filtered = true;
    }
    p.out.filtered[s] = filtered ? 1 : 0;
    p.out.n_used[s] = n_used;
    if (filtered) return 0;
#pragma unroll
    for (int i = 0; i < 10; ++i) mom.xtx[i] = 0.0;
#pragma unroll
    for (int i = 0; i < 4; ++i) mom.xty[i] = 0.0;
    mom.n_used = n_used;
    mom.p = 2;
    mom.state = 1;
    mom.pad = 0;
    const double su = p.sy_all - t.su_un, suu = p.syy_all - t.suu_un;
    mom.xtx[0] = double(n_used);
    mom.xtx[1] = colsum0;
    mom.xtx[2] =
        double(t.n_full0) + 0.25 * double(t.n_half) + (1.0 / 9.0) * double(t.n_third) + (4.0 / 9.0) * double(t.n_tthird);
    mom.xty[0] = su;
    mom.xty[1] = t.xty1;
    mom.sy = su;
    mom.syy = suu;
    return 1;
}

template <int WPT>
struct WordVec;
template <>
struct WordVec<1> {
    uint32_t v[1];
    __device__ __forceinline__ void load(const uint8_t* p) { v[0] = *reinterpret_cast<const uint32_t*>(p); }
};
template <>
struct WordVec<2> {
    uint32_t v[2];
    __device__ __forceinline__ void load(const uint8_t* p) {
        const uint2 t = *reinterpret_cast<const uint2*>(p);
        v[0] = t.x; v[1] = t.y;
    }
};
template <>
struct WordVec<4> {
    uint32_t v[4];
    __device__ __forceinline__ void load(const uint8_t* p) {
        const uint4 t = *reinterpret_cast<const uint4*>(p);
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    }
};

// Weight 1 + d/2 of a sample (d = its path-0 dosage in {0, 1/2, 1}) as a double: 1.0, 1.25, 1.5 are LINEAR in the 2-bit
// field 2d placed at mantissa bits 50, 51, so with the field (masked, at bit B >= 19 of a 32-bit word) the whole 64-bit
// pattern is ONE wide multiply-add, field * 2^(50 - B) + bits(1.0) (SASS IMAD.WIDE.U32 writes the register pair; no
// separate zeroing of the low word).  sum d*y = 2 (sum (1 + d/2) y - sum y) then needs one LOP3, one IMAD.WIDE and one
// DFMA per sample (nvcc turns a conditional add into two FSELs + DADD per class instead).
template <int B>
__device__ __forceinline__ double dosage_weight_at(uint32_t word, uint32_t lo_zero) {
    static_assert(B >= 18 && B <= 30, "field must sit at bits 18..30");
    // hi word = (field >> (B - 18)) + hi(1.0); the low word is a zero the compiler cannot see through (a kernel
    // parameter), one per accumulator: with a shared literal 0 ptxas funnels every weight through ONE register pair and
    // the 16 DFMAs of a word become a single dependent chain.
    const uint32_t hi = ((word & (3u << B)) >> (B - 18)) + 0x3FF00000u;
    return __hiloint2double(int(hi), int(lo_zero));
}
// 16 samples of one word: fields at bits 20..30 of w2 = fld (samples 10..15), of w1 = fld << 12 (samples 4..9) and at
// bits 24..30 of w0 = fld << 24 (samples 0..3)
#define STOAT_W(I, fld, w1, w0, z)                                                                            \
    ((I) >= 10 ? dosage_weight_at<((I) >= 10 ? 2 * (I) : 20)>(fld, z)                                         \
               : (I) >= 4 ? dosage_weight_at<((I) >= 4 && (I) < 10 ? 2 * (I) + 12 : 20)>(w1, z)               \
                          : dosage_weight_at<((I) < 4 ? 2 * (I) + 24 : 20)>(w0, z))

constexpr int kRegMaxWarps = 8;  // warps of a K3R CTA (256 threads)

// One path's presence words for this thread (codes as in quant_lean.cuh; branches are CTA-uniform).
template <int WPT>
__device__ __forceinline__ void reg_path_words(const uint8_t* rows_s, const uint8_t* locals, uint32_t code,
                                               uint32_t slot_bytes, uint32_t byte_off, uint32_t first_word,
                                               uint32_t n_hap, uint32_t (&out)[WPT]) {
    if (code == 0xFFu) {
#pragma unroll
        for (int k = 0; k < WPT; ++k) out[k] = 0u;
        return;
    }
    if (code == 0u) {
#pragma unroll
        for (int k = 0; k < WPT; ++k) out[k] = valid_word(first_word + k, n_hap);
        return;
    }
    WordVec<WPT> a;
    a.load(rows_s + uint32_t(locals[0]) * slot_bytes + byte_off);
#pragma unroll
    for (int k = 0; k < WPT; ++k) out[k] = a.v[k];
#pragma unroll
    for (uint32_t r = 1; r < 4; ++r)
        if (r < code) {  // code <= 4 on this path
            a.load(rows_s + uint32_t(locals[r]) * slot_bytes + byte_off);
#pragma unroll
            for (int k = 0; k < WPT; ++k) out[k] &= a.v[k];
        }
}


constexpr int kRegBufs = 4;  // partial-sum buffers: warps may run up to 4 snarls ahead of the slowest finisher
constexpr int kRegLag = 2;   // a stage is refilled two iterations after its snarl (nobody waits for the slowest reader)

// =================================================================================================
// K3R kernel.  grid = (CTAs per slice, slices); CTA b of a slice takes snarls b, b + gridDim.x, ...
// dynamic shared memory: [mbarriers: full[ring] empty[ring] ready[4] free[4] | pad to 128][ring stages of
// (64 B program + stage_rows slots)].  No CTA-wide barrier in the loop: warps run freely, coupled only through
//   full[st]   TMA landed (tx count)                      empty[st]  every warp has read its words of the stage
//   free[b]    the last warp to publish its partial sums in buffer b has added them up (smem counter arrived[b])
// =================================================================================================
template <int WPT>
__global__ void __launch_bounds__(kRegMaxWarps * 32) dosage_moments_reg_kernel(const K3RParams q) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ uint32_t red_i[kRegBufs][4][kRegMaxWarps];
    __shared__ double red_d[kRegBufs][3][kRegMaxWarps];
    __shared__ uint32_t arrived[kRegBufs];
    const K3Params& p = q.k;
    const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, n_warps = blockDim.x >> 5;
    const uint32_t n_words = p.pitch / 4;
    const uint32_t slice = blockIdx.y, slice_first = slice * q.slice_words;
    const uint32_t slice_len = min(q.slice_words, n_words - slice_first);  // words of this slice (multiple of 4)
    const uint32_t slot_bytes = q.slice_words * 4, slice_bytes = slice_len * 4;
    const uint32_t stage_bytes = kStageBlobBytes + p.stage_rows * slot_bytes;
    const uint32_t ring = q.n_ring;
    if (tid < kRegBufs) arrived[tid] = 0u;
    uint64_t* full = reinterpret_cast<uint64_t*>(smem);
    uint64_t* empty = full + ring;
    uint64_t* ready = empty + ring;
    uint64_t* bfree = ready + kRegBufs;
    uint8_t* stages = smem + (((2 * ring + 2 * kRegBufs) * 8 + 127) & ~127u);

    if (tid == 0) {
        for (uint32_t i = 0; i < ring; ++i) {
            mbar_init(&full[i], 1);
            mbar_init(&empty[i], n_warps);
        }
        for (uint32_t i = 0; i < kRegBufs; ++i) {
            mbar_init(&ready[i], n_warps);
            mbar_init(&bfree[i], 1);
        }
        mbar_fence_init();
    }
    for (uint32_t i = tid; i < kRegBufs * 4 * kRegMaxWarps; i += blockDim.x) (&red_i[0][0][0])[i] = 0u;
    for (uint32_t i = tid; i < kRegBufs * 3 * kRegMaxWarps; i += blockDim.x) (&red_d[0][0][0])[i] = 0.0;
    // this thread's samples: words [first_word, first_word + WPT) of the row, 16 samples each
    const uint32_t local_word = tid * WPT, first_word = slice_first + local_word;
    const bool active = local_word < slice_len;
    const uint32_t byte_off = active ? local_word * 4 : 0u;  // idle threads read word 0 of the slot and discard it
    double y[16 * WPT];
#pragma unroll
    for (int k = 0; k < WPT; ++k)
#pragma unroll
        for (int i = 0; i < 16; ++i) y[k * 16 + i] = active ? __ldg(p.ys + size_t(i) * n_words + first_word + k) : 0.0;
    double y_total;  // in the order the accumulators see the samples
    {
        double t0 = 0.0, t1 = 0.0, t2 = 0.0, t3 = 0.0;
#pragma unroll
        for (int k = 0; k < WPT; ++k)
#pragma unroll
            for (int i = 0; i < 16; i += 4) {
                t0 += y[k * 16 + i];
                t1 += y[k * 16 + i + 1];
                t2 += y[k * 16 + i + 2];
                t3 += y[k * 16 + i + 3];
            }
        y_total = (t0 + t1) + (t2 + t3);
    }
    uint32_t valid[WPT];
#pragma unroll
    for (int k = 0; k < WPT; ++k) valid[k] = active ? (valid_word(first_word + k, p.n_hap) & 0x55555555u) : 0u;
    const uint32_t live = active ? 0xFFFFFFFFu : 0u;
    __syncthreads();

    const uint64_t stride = gridDim.x;
    const uint64_t policy = make_evict_first_policy();
    if (warp == 0) {
        uint64_t s = blockIdx.x;
        for (uint32_t st = 0; st < ring && s < p.n_snarls; ++st, s += stride)
            issue_stage_slice(load_rec_word(p.rec, s, lane), lane, stages + size_t(st) * stage_bytes, &full[st], p.rows,
                              p.blob, p.pitch, slice_first * 4, slice_bytes, slot_bytes, policy);
    }

    uint32_t stage = 0, parity = 0;      // stage / phase of the current snarl
    uint32_t rstage = 0, rparity = 0;    // stage / phase of the snarl kRegLag iterations back (warp 0 refills it)
    uint32_t buf = 0, bphase = 0, it = 0;
    uint32_t rec_ahead = 0;  // warp 0: record words of the snarl the NEXT iteration stages
    if (warp == 0) {
        const uint64_t s0 = uint64_t(blockIdx.x) + uint64_t(ring - kRegLag) * stride;
        if (s0 < p.n_snarls) rec_ahead = load_rec_word(p.rec, s0, lane);
    }
    for (uint64_t s = blockIdx.x; s < p.n_snarls; s += stride, ++it) {
        // ---- warp 0: refill the stage every warp left kRegLag iterations ago ------------------------------
        if (warp == 0) {
            // (the record of the snarl to stage was fetched one iteration earlier: its latency is off this warp's path)
            const uint64_t s_new = s + uint64_t(ring - kRegLag) * stride;
            const uint32_t w = rec_ahead;
            rec_ahead = (s_new + stride < p.n_snarls) ? load_rec_word(p.rec, s_new + stride, lane) : 0u;
            if (it >= uint32_t(kRegLag)) {
                if (s_new < p.n_snarls) {
                    mbar_wait(&empty[rstage], rparity);
                    issue_stage_slice(w, lane, stages + size_t(rstage) * stage_bytes, &full[rstage], p.rows, p.blob,
                                      p.pitch, slice_first * 4, slice_bytes, slot_bytes, policy);
                }
                if (++rstage == ring) {
                    rstage = 0;
                    rparity ^= 1u;
                }
            }
        }
        mbar_wait(&full[stage], parity);
        const uint8_t* blob = stages + size_t(stage) * stage_bytes;
        const uint8_t* rows_s = blob + kStageBlobBytes;
        const uint4 hdr = *reinterpret_cast<const uint4*>(blob);  // np, shape class | p0 | codes, locals ...
        const uint32_t np = hdr.x & 0xFFu, shape = (hdr.x >> 8) & 0xFFu;
        const uint32_t code0 = hdr.z & 0xFFu, code1 = (hdr.z >> 8) & 0xFFu;
        const bool lean = np == 2 && (code0 == 0xFFu || code0 <= 4u) && (code1 == 0xFFu || code1 <= 4u);
        uint32_t a[WPT], b[WPT];
        if (shape == 1u) {  // two paths, two checked edges each: locals in bytes 10..13
            WordVec<WPT> r0, r1, r2, r3;
            r0.load(rows_s + ((hdr.z >> 16) & 0xFFu) * slot_bytes + byte_off);
            r1.load(rows_s + (hdr.z >> 24) * slot_bytes + byte_off);
            r2.load(rows_s + (hdr.w & 0xFFu) * slot_bytes + byte_off);
            r3.load(rows_s + ((hdr.w >> 8) & 0xFFu) * slot_bytes + byte_off);
#pragma unroll
            for (int k = 0; k < WPT; ++k) {
                a[k] = r0.v[k] & r1.v[k] & live;
                b[k] = r2.v[k] & r3.v[k] & live;
            }
        } else if (lean) {
            const uint8_t* loc0 = blob + kBlobHeaderQ + 2;
            const uint8_t* loc1 = loc0 + ((code0 == 0xFFu) ? 0u : code0);
            reg_path_words<WPT>(rows_s, loc0, code0, slot_bytes, byte_off, first_word, p.n_hap, a);
            reg_path_words<WPT>(rows_s, loc1, code1, slot_bytes, byte_off, first_word, p.n_hap, b);
#pragma unroll
            for (int k = 0; k < WPT; ++k) {
                a[k] &= live;
                b[k] &= live;
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[stage]);  // this warp is done with the stage

        uint32_t cA = 0, cB = 0, cC = 0, cD = 0;  // al0 | al1 << 16, n_used | n_full0 << 16, n_half, n_third | n_tthird << 16
        double xty = 0.0, su_un = 0.0, suu_un = 0.0;
        if (lean) {
            double acc0 = 0.0, acc1 = 0.0, acc2 = 0.0, acc3 = 0.0;
            const uint32_t z0 = q.zero4[0], z1 = q.zero4[1], z2 = q.zero4[2], z3 = q.zero4[3];
            uint32_t rare = 0;
#pragma unroll
            for (int k = 0; k < WPT; ++k) {
                const LeanMasks m = lean_masks(a[k], b[k]);
                cA += __popc(a[k]) | (__popc(b[k]) << 16);
                cB += __popc(m.used) | (__popc(m.full0) << 16);
                cC += __popc(m.half);
                rare |= (~m.used & valid[k]) | m.third | m.tthird;
                const uint32_t fld = m.half | (m.full0 << 1);  // 2-bit field per sample: 2 * dosage
                const uint32_t w1 = fld << 12, w0 = fld << 24;
                const double* yk = y + k * 16;
                acc0 = fma(STOAT_W(0, fld, w1, w0, z0), yk[0], acc0);
                acc1 = fma(STOAT_W(1, fld, w1, w0, z1), yk[1], acc1);
                acc2 = fma(STOAT_W(2, fld, w1, w0, z2), yk[2], acc2);
                acc3 = fma(STOAT_W(3, fld, w1, w0, z3), yk[3], acc3);
                acc0 = fma(STOAT_W(4, fld, w1, w0, z0), yk[4], acc0);
                acc1 = fma(STOAT_W(5, fld, w1, w0, z1), yk[5], acc1);
                acc2 = fma(STOAT_W(6, fld, w1, w0, z2), yk[6], acc2);
                acc3 = fma(STOAT_W(7, fld, w1, w0, z3), yk[7], acc3);
                acc0 = fma(STOAT_W(8, fld, w1, w0, z0), yk[8], acc0);
                acc1 = fma(STOAT_W(9, fld, w1, w0, z1), yk[9], acc1);
                acc2 = fma(STOAT_W(10, fld, w1, w0, z2), yk[10], acc2);
                acc3 = fma(STOAT_W(11, fld, w1, w0, z3), yk[11], acc3);
                acc0 = fma(STOAT_W(12, fld, w1, w0, z0), yk[12], acc0);
                acc1 = fma(STOAT_W(13, fld, w1, w0, z1), yk[13], acc1);
                acc2 = fma(STOAT_W(14, fld, w1, w0, z2), yk[14], acc2);
                acc3 = fma(STOAT_W(15, fld, w1, w0, z3), yk[15], acc3);
            }
            xty = 2.0 * (((acc0 + acc1) + (acc2 + acc3)) - y_total);  // 2 (sum (1 + d/2) y - sum y)
            if (rare) {  // samples on neither path / on three path-haplotypes (dosages 1/3, 2/3): a few lanes, rarely
#pragma unroll
                for (int k = 0; k < WPT; ++k) {
                    const LeanMasks m = lean_masks(a[k], b[k]);
                    const uint32_t un = ~m.used & valid[k];
                    cD += __popc(m.third) | (__popc(m.tthird) << 16);
                    double s3 = 0.0, s23 = 0.0;
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        const double yy = y[k * 16 + i];
                        if (un & (1u << (2 * i))) {
                            su_un += yy;
                            suu_un = fma(yy, yy, suu_un);
                        }
                        if (m.third & (1u << (2 * i))) s3 += yy;
                        if (m.tthird & (1u << (2 * i))) s23 += yy;
                    }
                    xty += (1.0 / 3.0) * s3 + (2.0 / 3.0) * s23;
                }
            }
            // ---- warp level: packed integer reductions, FP64 by shuffles (the rare sums only when some lane has any)
            cA = __reduce_add_sync(0xffffffffu, cA);
            cB = __reduce_add_sync(0xffffffffu, cB);
            cC = __reduce_add_sync(0xffffffffu, cC);
            xty = warp_sum(xty);
            if (__any_sync(0xffffffffu, rare != 0u)) {
                su_un = warp_sum(su_un);
                suu_un = warp_sum(suu_un);
                cD = __reduce_add_sync(0xffffffffu, cD);
            }
        }
        // ---- publish this warp's partial sums; the LAST warp to arrive adds them up ---------------------------
        // (a designated finishing warp would have to wait for the slowest warp and then delays its own next snarl: with a
        //  rotating finisher that wait becomes a serial chain through the whole CTA)
        if (it >= uint32_t(kRegBufs)) mbar_wait(&bfree[buf], bphase ^ 1u);  // the buffer's previous snarl is finished
        uint32_t last = 0;
        if (lane == 0) {
            red_i[buf][0][warp] = cA;
            red_i[buf][1][warp] = cB;
            red_i[buf][2][warp] = cC;
            red_i[buf][3][warp] = cD;
            red_d[buf][0][warp] = xty;
            red_d[buf][1][warp] = su_un;
            red_d[buf][2][warp] = suu_un;
            __threadfence_block();
            last = (atomicAdd(&arrived[buf], 1u) == n_warps - 1u) ? 1u : 0u;
        }
        last = __shfl_sync(0xffffffffu, last, 0);
        if (last) {
            __threadfence_block();
            if (lean) {
                // lanes 0..6: one class count each; lanes 8..10: one FP64 sum each; added in warp order
                if (lane < 7) {
                    const uint32_t src = (lane < 2) ? 0u : (lane < 4) ? 1u : (lane == 4) ? 2u : 3u;
                    const uint32_t sh = (lane == 1 || lane == 3 || lane == 6) ? 16u : 0u;
                    const uint32_t mask = (lane == 4) ? 0xFFFFFFFFu : 0xFFFFu;
                    uint32_t v = 0;
#pragma unroll
                    for (int w = 0; w < kRegMaxWarps; ++w) v += (red_i[buf][src][w] >> sh) & mask;
                    q.part.i[(size_t(slice) * 7 + lane) * q.part.stride + s] = v;
                } else if (lane >= 8 && lane < 11) {
                    double v = 0.0;
#pragma unroll
                    for (int w = 0; w < kRegMaxWarps; ++w) v += red_d[buf][lane - 8][w];
                    q.part.d[(size_t(slice) * 3 + (lane - 8)) * q.part.stride + s] = v;
                } else if (lane == 11 && slice == 0) {
                    p.mom[s].state = 2;  // K4 builds the moments from the partial sums
                }
            } else if (lane == 0 && slice == 0) {
                p.mom[s].state = 0;
                if (np) {  // np == 0: prep already put the snarl on the general list
                    const uint32_t slot = atomicAdd(q.slow_count, 1u);
                    q.slow_list[slot] = uint32_t(s);
                }
            }
            __syncwarp();
            if (lane == 0) {
                arrived[buf] = 0u;
                mbar_arrive(&bfree[buf]);  // (release: the reset is visible before anybody passes the wait above)
            }
        }
        if (++stage == ring) {
            stage = 0;
            parity ^= 1u;
        }
        if (++buf == uint32_t(kRegBufs)) {
            buf = 0;
            bphase ^= 1u;
        }
    }
}
