"""bench_eqtl.py — the dense eQTL scan leg of bench.py (`python bench.py --trait eqtl`), BASELINE.json configs[3] ("C4"):
synthetic 200,000 biallelic snarls x 5,000 samples x 20,000 gene phenotypes, every (snarl, gene) pair one OLS
(EQTLSnarlAnalyzer, src/snarl_analyzer.cpp:493-539).  A step = one full scan over the resident matrix / expression.

  value      (snarl, gene) OLS tests / s, whole job, inputs resident in HBM
  e2e        same through stoat_gpu_eqtl_dense_summary with HOST buffers (matrix + expression H2D, summary D2H)
  roofline   E1 eqtl_dense_dmma: 2 * n * S_dense * G flops / its CUDA-event time, against an FP64 GEMM measured live
             (MEASURED_PEAKS.json has no FP64 figure): torch.matmul f64 8192^3 = cuBLAS DGEMM, best of 5
  cpu_baseline / --impl reference   the reference's own EQTLSnarlAnalyzer (oracle/_ref) or the oracle port on a sample
"""
from __future__ import annotations

import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
METRIC = "snarl-gene OLS tests/sec (dense eQTL scan)"
UNIT = "pair-tests/s"
BIALLELIC = dict(frac_three=0.0, frac_node0=0.0)  # C4 is "biallelic snarls" 


def config_dict(args, n_gpus):
    return {
        "workload": f"C4 synthetic {args.snarls} biallelic snarls x {args.samples} samples x {args.genes} genes, "
                    f"dense eQTL scan (one OLS per pair), per GPU",
        "snarls_per_gpu": args.snarls, "samples": args.samples, "genes": args.genes,
        "sharding": ("single GPU" if n_gpus <= 1 else
                     f"gene-sharded: {n_gpus} ranks scan the same snarls against {args.genes} / {n_gpus} genes each, per-snarl "
                     f"summaries merged by one all-reduce per step" if getattr(args, "sharding", "snarl") == "gene" else
                     "snarl-sharded, no collective"),
        "l2": "the operand images (expression limbs 0.8 GB, dosage 0.9 GB) exceed the 126 MB L2; a gene group (~10 MB) and a "
              "panel of slot tiles are L2 resident by design, every step streams the rest from HBM, no flush needed",
        "p_threshold": args.p_threshold, "thresholds": {"min_individuals": 3, "min_haplotypes": 5, "maf": 0.05},
        "covariates": (f"{args.ncov} covariates: every snarl on the general kernel (K3G); the tensor-core kernel the roofline "
                       "object names does not run" if getattr(args, "ncov", 0) else "none"),
    }


def expression_numpy(n_genes, n_samples, seed=7):
    """e_g * (1 + U(-1, 1)), e_g ~ U(0, 8): tests/simu/scripts/simulate_test_dataset.py:75-88."""
    rng = np.random.default_rng(seed)
    return rng.uniform(0, 8, size=(n_genes, 1)) * (1.0 + rng.uniform(-1, 1, size=(n_genes, n_samples)))


# ---------------------------------------------------------------------------------------------------
# CPU arm
# ---------------------------------------------------------------------------------------------------
def cpu_kind():
    try:
        from oracle import ref

        if os.path.exists(ref.ref_build.LIB) or ref.available():
            ref.lib()
            return "reference"
    except Exception:
        pass
    return "port"


def cpu_run(d, expr, threads, kind):
    """One pass of the CPU eQTL loop over the sample: every gene is in every snarl's window.  -> (seconds, pairs)."""
    import oracle

    S, G = d["paths"].n_snarls, expr.shape[0]
    if kind == "reference":
        from oracle import ref

        nodes = d.get("_nodes")
        if nodes is None:
            nodes = d["_nodes"] = ref.nodes_from_structure(d["structure"])
        r = ref.run("eqtl", d["bits"], d["n_rows"], d["n_cols"], d["paths"].snarl_path_off, nodes, expression=expr,
                    window=1 << 40, n_threads=threads, tsv_path=os.path.join("/tmp", f"stoat_cpu_eqtl_{os.getpid()}.tsv"),
                    parse=False)
        return r["seconds"], S * G
    P = oracle.Paths(d["paths"].snarl_path_off, d["paths"].path_edge_off, d["paths"].edge_row)
    r = oracle.eqtl_batch(d["bits"], d["n_cols"], expr, P, n_threads=threads)
    return r["seconds"], S * G


CPU_WHAT = {"reference": "the reference's own EQTLSnarlAnalyzer sources (oracle/_ref: table once per snarl, Eigen OLS per gene, "
                         "writer; Boost tails from stand-ins) incl. per-row TSV write",
            "port": "oracle port of the reference's eQTL loop"}


def cpu_sample(args, target_seconds, n_threads):
    """(dataset, expression, threads, kind) sized for about target_seconds of CPU work; n_threads = the host threads this
    process may use (bench.host_threads(): CPU affinity and cgroup quota, not OMP_NUM_THREADS)."""
    from stoat_b200 import synth

    kind = cpu_kind()
    probe = synth.numpy_dataset(8, args.samples, seed=42, **BIALLELIC)
    pe = expression_numpy(16, args.samples)
    best = (1, 0.0)
    for th in sorted({1, n_threads}):
        secs, pairs = cpu_run(probe, pe, th, kind)
        if pairs / max(secs, 1e-9) > best[1]:
            best = (th, pairs / max(secs, 1e-9))
    threads, rate = best
    want = max(64, int(rate * target_seconds))
    n_genes = int(min(args.genes, max(16, round(want ** 0.5))))
    n_snarls = int(min(2000, max(8, want // n_genes)))
    d = synth.numpy_dataset(n_snarls, args.samples, seed=42, **BIALLELIC)
    return d, expression_numpy(n_genes, args.samples), threads, kind


def cpu_baseline(args, target_seconds, n_threads):
    d, expr, threads, kind = cpu_sample(args, target_seconds, n_threads)
    secs, pairs = cpu_run(d, expr, threads, kind)
    return {"value": pairs / secs, "unit": UNIT, "cores": threads, "kind": kind,
            "sample": f"{d['paths'].n_snarls} snarls x {expr.shape[0]} genes x {args.samples} samples ({pairs} pairs) of the same "
                      f"generators, {CPU_WHAT[kind]}, {secs:.2f} s"}


def reference_line(args, n_threads):
    import oracle

    oracle.build()
    d, expr, threads, kind = cpu_sample(args, 4.0, n_threads)
    for _ in range(args.warmup):
        cpu_run(d, expr, threads, kind)
    total, pairs = 0.0, 0
    for _ in range(args.steps):
        s, p = cpu_run(d, expr, threads, kind)
        total += s
        pairs += p
    value = pairs / total
    sample = (f"{d['paths'].n_snarls} snarls x {expr.shape[0]} genes x {args.samples} samples per step, same generators, "
              f"{CPU_WHAT[kind]}")
    return {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": config_dict(args, args.gpus),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}


def run_reference_arm(args, n_threads):
    if int(os.environ.get("RANK", "0")) != 0:
        return
    print(json.dumps(reference_line(args, n_threads)))


# ---------------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------------
def dgemm_peak_tflops(torch, dev, n=8192, reps=5):
    """cuBLAS DGEMM through torch.matmul, CUDA events, best of `reps` (the FP64 tensor-core roofline of this box)."""
    a = torch.randn((n, n), device=dev, dtype=torch.float64)
    b = torch.randn((n, n), device=dev, dtype=torch.float64)
    for _ in range(2):
        a @ b
    torch.cuda.synchronize()
    best = float("inf")
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        a @ b
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    del a, b
    return 2.0 * n ** 3 / (best * 1e-3) / 1e12


def int8_peak_tops(torch, dev, n=8192, reps=5):
    """int8 x int8 -> int32 GEMM through torch._int_mm (cuBLASLt), CUDA events, best of `reps`: the int8 tensor-core roofline
    of this box in TOP/s (2 ops per multiply-add).  None when the build has no int8 matmul."""
    try:
        a = torch.randint(-64, 64, (n, n), device=dev, dtype=torch.int8)
        b = torch.randint(-64, 64, (n, n), device=dev, dtype=torch.int8)
        for _ in range(2):
            torch._int_mm(a, b)
        torch.cuda.synchronize()
        best = float("inf")
        for _ in range(reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            torch._int_mm(a, b)
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        del a, b
        return 2.0 * n ** 3 / (best * 1e-3) / 1e12
    except Exception:
        return None


def parity_spot_check(ctx_summary, d, expr_dev, args, n_check, n_threads, gene0=0, covar=None):
    """`n_check` snarls SAMPLED at a fixed stride over the timed scan x every gene through the CPU oracle (one OLS per pair,
    as the reference: n_check x G = 10^6 rows at C4), against the device's per-snarl summary: min p (1e-9
    relative), its gene, the count below the threshold."""
    import time

    import oracle

    st = d["structure"]
    S = d["paths"].n_snarls
    n = min(n_check, S)
    sel = np.unique(np.linspace(0, S - 1, n).astype(np.int64))
    n = sel.size
    row_bytes = d["n_cols"] // 8
    p = d["paths"]
    spo_all, peo_all = np.asarray(p.snarl_path_off, np.int64), np.asarray(p.path_edge_off, np.int64)
    torch = __import__("torch")
    # the sampled snarls' rows (private and contiguous per snarl in this generator), compacted; edge rows renumbered
    row_idx, spo, peo, er, base = [], [0], [0], [], 0
    for s_ in sel:
        r0, nr = int(st.row0[s_]), int(st.n_rows_of[s_])
        row_idx.append(np.arange(r0, r0 + nr))
        for pth in range(spo_all[s_], spo_all[s_ + 1]):
            rows = np.asarray(p.edge_row[peo_all[pth]:peo_all[pth + 1]], np.int64)
            er.extend(np.where(rows == 0xFFFFFFFF, 0xFFFFFFFF, rows - r0 + base).tolist())
            peo.append(len(er))
        spo.append(len(peo) - 1)
        base += nr
    row_idx = torch.as_tensor(np.concatenate(row_idx), device=d["bits_dev"].device)
    bits = d["bits_dev"].view(-1, row_bytes)[row_idx].contiguous().cpu().numpy().reshape(-1)
    expr = expr_dev.cpu().numpy()
    sub = oracle.Paths(np.array(spo, np.uint64), np.array(peo, np.uint64), np.array(er, np.uint32))
    ctx_summary = {k: np.asarray(v)[sel] for k, v in ctx_summary.items()}
    t0 = time.time()
    ref = oracle.eqtl_batch(bits, d["n_cols"], expr, sub, covar=covar, n_threads=n_threads)
    G, thr, rtol = expr.shape[0], args.p_threshold, 1e-9
    pv = ref["p_value"].reshape(n, G)
    ok = ref["code"].reshape(n, G) == 0
    bad = []
    for s in range(n):
        mp, ag, nb = ctx_summary["min_p"][s], int(ctx_summary["argmin_gene"][s]) - gene0, int(ctx_summary["n_below"][s])
        if ref["filtered"][s] or not ok[s].any():
            if not (np.isnan(mp) and nb == 0):
                bad.append((s, "filtered"))
            continue
        ps = np.where(ok[s], pv[s], np.inf)
        g = int(np.argmin(ps))
        near = np.abs(ps - thr) <= rtol * thr
        lo, hi = int(((ps < thr) & ~near).sum()), int(((ps < thr) | near).sum())
        if not (abs(mp - ps[g]) <= rtol * ps[g] + 1e-300):
            bad.append((s, "min_p", float(mp), float(ps[g])))
        elif ag != g and not abs(ps[ag] - ps[g]) <= rtol * ps[g]:
            bad.append((s, "argmin", ag, g))
        elif not lo <= nb <= hi:
            bad.append((s, "n_below", nb, lo, hi))
    out = {"checked_snarls": int(n), "checked_pair_tests": int(n * G), "mismatches": len(bad),
           "sampling": "snarls at a fixed stride over the whole timed batch, every gene",
           "against": "oracle port (oracle/stoat_oracle.cpp), one OLS per (snarl, gene) on the timed matrix's own rows and "
                      "the timed expression matrix; min p 1e-9 relative, its gene, count below the threshold",
           "seconds": round(time.time() - t0, 2)}
    if bad:
        out["first_mismatch"] = str(bad[0])
    return out


def measure(args, env, ClockSampler, n_threads, with_clocks=True):
    """The C4 record on this rank's GPU; `env` = bench.Env (torch, process group).  -> the JSON line's dict on rank 0."""
    import stoat_b200
    from stoat_b200 import synth

    torch, dist = env.torch, env.dist
    world, rank, local, dev = env.world, env.rank, env.local, env.dev
    barrier = env.barrier

    S, n, G = args.snarls, args.samples, args.genes
    # snarl-sharded (default): every rank its own snarl shard against every gene (weak scaling, no exchange).
    # gene-sharded (--sharding gene, SURVEY §8e "split genes (columns of Y) across GPUs"): every rank the SAME snarls
    # against G / world genes (strong scaling); the per-snarl summaries are merged with one small all-reduce per step.
    gene_sharded = getattr(args, "sharding", "snarl") == "gene" and world > 1
    d = synth.torch_dataset(S, n, dev, seed=42 + (0 if gene_sharded else rank), **BIALLELIC)
    gen = torch.Generator(device=dev)
    gen.manual_seed(7)
    base = torch.rand((G, 1), device=dev, generator=gen, dtype=torch.float64) * 8
    expr = base * (1 + (torch.rand((G, n), device=dev, generator=gen, dtype=torch.float64) * 2 - 1))
    gene0 = 0
    if gene_sharded:
        if G % world:
            raise SystemExit("bench.py --trait eqtl --sharding gene: the gene count must be a multiple of the GPU count")
        per = (G + world - 1) // world
        gene0 = min(G, rank * per)
        expr = expr[gene0:min(G, gene0 + per)].contiguous()
    G_local = expr.shape[0]
    # 10 % of the genes depend on a snarl's dosage (simulate_test_dataset.py:75-88): add the alt-path dosage of snarl g
    peak = dgemm_peak_tflops(torch, dev) if rank == 0 else None
    umma = os.environ.get("STOAT_GPU_EQTL", "") not in ("dmma", "general")
    peak_i8 = int8_peak_tops(torch, dev) if (rank == 0 and umma) else None
    ctx = stoat_b200.Context(local)
    ctx.set_matrix_dev(d["bits_dev"].data_ptr(), d["n_rows"], d["n_cols"])
    ctx.stage_paths(d["paths"])
    ctx.set_expression_dev(expr.data_ptr(), G_local)
    # SURVEY §8d "no covariates, then ncov = 3": with covariates every snarl takes the general kernel (K3G: the table and
    # (X'X)^-1 once per snarl, then two sweeps over the samples per gene) - there is no tensor-core form of it, see DESIGN §9
    covar = None
    if getattr(args, "ncov", 0):
        covar = np.random.default_rng(11).normal(size=(n, args.ncov))
    ctx.set_covariates(covar)
    stream = torch.cuda.ExternalStream(ctx.stream, device=dev)
    merged = {}

    def run_scan():
        ctx.eqtl_dense_run_staged(p_threshold=args.p_threshold)
        if gene_sharded:
            # merge of the per-snarl summaries: min p (and whose gene it is) and the count, S x 16 bytes per rank
            sm = ctx.eqtl_summary_fetch()
            mp = torch.as_tensor(np.nan_to_num(sm["min_p"], nan=np.inf), device=dev)
            ag = torch.as_tensor(sm["argmin_gene"].astype(np.int64) + gene0, device=dev)
            nb = torch.as_tensor(sm["n_below"].astype(np.int64), device=dev)
            best = mp.clone()
            dist.all_reduce(best, op=dist.ReduceOp.MIN)
            cand = torch.where(mp == best, ag, torch.full_like(ag, 1 << 40))
            dist.all_reduce(cand, op=dist.ReduceOp.MIN)  # ties: the smallest gene index, as a single scan would report
            dist.all_reduce(nb, op=dist.ReduceOp.SUM)
            merged["min_p"] = torch.where(torch.isinf(best), torch.full_like(best, float("nan")), best).cpu().numpy()
            merged["argmin_gene"] = cand.cpu().numpy()
            merged["n_below"] = nb.cpu().numpy()
    for _ in range(max(args.warmup, 3)):
        run_scan()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0 and with_clocks:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e1_ms, rest_ms, launches, n_general = [], [], 0, 0
    e0.record(stream)
    for _ in range(args.steps):
        run_scan()
        t = ctx.timing()
        e1_ms.append(t["table_ms"])
        rest_ms.append(t["stats_ms"])
        launches += t["launches"]
        n_general = t["n_general"]
    e1.record(stream)
    barrier()
    clocks = sampler.stop() if (rank == 0 and with_clocks) else None
    tmax = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    lsum = torch.tensor([launches], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        dist.all_reduce(lsum, op=dist.ReduceOp.SUM)
    elapsed_ms = float(tmax.item())
    value = (1 if gene_sharded else world) * S * G * args.steps / (elapsed_ms * 1e-3)
    summary = ctx.eqtl_summary_fetch()
    n_dense = int(S - n_general - int(np.isnan(summary["min_p"]).sum()))
    parity = None
    if not getattr(args, "no_parity", False):
        if gene_sharded:
            # the MERGED summary against the oracle over every gene: gather the ranks' expression slices on rank 0
            parts = [torch.empty_like(expr) for _ in range(world)]
            dist.all_gather(parts, expr)
            if rank == 0:
                parity = parity_spot_check(merged, d, torch.cat(parts, dim=0), args, 50, n_threads, covar=covar)
            del parts
        elif rank == 0:
            parity = parity_spot_check(summary, d, expr, args, 50, n_threads, covar=covar)

    e2e = None
    if not args.no_e2e and not gene_sharded:
        host_bits = torch.empty(d["bits_dev"].numel(), dtype=torch.uint8, pin_memory=True)
        host_bits.copy_(d["bits_dev"])
        host_expr = torch.empty((G, n), dtype=torch.float64, pin_memory=True)
        host_expr.copy_(expr)
        torch.cuda.synchronize()
        expr_np = host_expr.numpy()

        def e2e_step():
            ctx.set_matrix_ptr(host_bits.data_ptr(), d["n_rows"], d["n_cols"])
            return ctx.eqtl_dense_summary(d["paths"], expr_np, p_threshold=args.p_threshold)

        e2e_step()
        barrier()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record(stream)
        for _ in range(args.e2e_steps):
            out = e2e_step()
        f1.record(stream)
        barrier()
        te = torch.tensor([f0.elapsed_time(f1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        p = d["paths"]
        h2d = host_bits.numel() + host_expr.numel() * 8 + p.snarl_path_off.nbytes + p.path_edge_off.nbytes + p.edge_row.nbytes
        d2h = sum(v.nbytes for v in out.values())
        e2e = {"value": world * S * G * args.e2e_steps / (float(te.item()) * 1e-3), "unit": UNIT,
               "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h), "ms_per_step": float(te.item()) / args.e2e_steps,
               "includes": "set_matrix (H2D + K0 repack) + stage (H2D CSR, prep) + expression H2D + centre + E0 + E1 + E2 "
                           "+ general-path rest + D2H of the per-snarl summary"}
        del host_bits, host_expr

    line = None
    if rank == 0:
        k_ms = float(np.mean(e1_ms))
        flops = 2.0 * n * n_dense * G_local
        if umma:
            # the exact int8 form: 2 d (int8) against the 8 signed 7-bit limbs of the centred expression, INT32 accumulators in
            # tensor memory.  Algorithmic ops = 8 limb columns per gene; the FP64-equivalent rate is what the FP64 form would need.
            ops = 8.0 * flops
            achieved = ops / (k_ms * 1e-3) / 1e12
            nominal = 4500.0  # dense int8 = dense fp8 on B200 (NVIDIA's 9 POPS figure is 2:4 sparse)
            pk = peak_i8 if peak_i8 else nominal
            roofline = {"bound": "tensor", "achieved": achieved, "peak": pk, "unit": "TFLOP/s", "frac": achieved / pk,
                        "traffic": None,
                        "kernel": "eqtl_dense_umma (EU1: tcgen05.mma kind::i8, M 128 x N 256 x K 32, accumulators in TMEM, per-pair "
                                  "OLS epilogue from tcgen05.ld) incl. the expansion of the dosage codes into the A operand",
                        "ops": "int8 multiply-adds x 2 (TOP/s reported under the contract's TFLOP/s unit)",
                        "algorithmic_flops_per_launch": ops, "kernel_ms": k_ms,
                        "peak_source": ("int8 GEMM measured live in this run: torch._int_mm 8192^3 (cuBLASLt), best of 5, CUDA "
                                        "events" if peak_i8 else "nominal dense int8 rate of B200 (torch._int_mm unavailable)") +
                                       "; MEASURED_PEAKS.json carries no int8 figure",
                        "fp64_equivalent": {"tflops": flops / (k_ms * 1e-3) / 1e12, "dgemm_peak_tflops": peak,
                                            "note": "2 n S G flops, what the FP64 (DMMA) form of the same contraction executes; "
                                                    "STOAT_GPU_EQTL=dmma runs that form (0.87 of cuBLAS DGEMM, 1.17 s per scan)"},
                        "limiter": "L2 -> SM operand traffic (48 KB per 4 MMAs: 171 int8 ops per byte against the ~600 the tensor "
                                   "rate would need), see DESIGN.md",
                        "snarls_on_tensor_path": n_dense, "snarls_on_general_path": int(n_general),
                        "other_kernels_ms": {"E0 dosage codes + E2 finish + general-path rest": float(np.mean(rest_ms))}}
        else:
            achieved = flops / (k_ms * 1e-3) / 1e12
            roofline = {"bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
                        "traffic": None, "kernel": "eqtl_dense_dmma (E1: DMMA.8x8x4 contraction + per-pair OLS epilogue)",
                        "algorithmic_flops_per_launch": flops, "kernel_ms": k_ms,
                        "peak_source": "FP64 GEMM measured live in this run: torch.matmul f64 8192^3 (cuBLAS DGEMM), best of 5, "
                                       "CUDA events; MEASURED_PEAKS.json carries no FP64 figure",
                        "snarls_on_tensor_path": n_dense, "snarls_on_general_path": int(n_general),
                        "other_kernels_ms": {"E0 dosage codes + E2 finish + general-path rest": float(np.mean(rest_ms))}}
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": elapsed_ms / args.steps, "higher_is_better": True,
                "scaling": "strong" if gene_sharded else "weak", "vs_baseline": None,
                "dtype": ("int8 x int8 -> int32 exact contraction (2-bit dosage codes x 8 limbs of the f64 expression, tcgen05), f64 "
                          "statistics" if umma else "f64 (2-bit dosage codes x f64 expression, DMMA)"),
                "data": "synthetic (seed 42, drawn on device)", "config": config_dict(args, world), "clocks": clocks,
                "e2e": e2e, "gpu_launches": int(lsum.item()), "roofline": roofline, "parity": parity}
        if world == 1 and not args.no_cpu_baseline:
            import oracle

            oracle.build()
            line["cpu_baseline"] = cpu_baseline(args, args.cpu_seconds, n_threads)
    ctx.close()
    del d, expr
    torch.cuda.empty_cache()
    return line if rank == 0 else None
