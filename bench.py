#!/usr/bin/env python3
"""bench.py — snarl-path tests/sec of the `stoat vcf` binary hot path (chi2 + Fisher) on B200.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
  torchrun --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...        (one rank per GPU)

Workload (BASELINE.json configs[2], "C3"): synthetic 1,000,000 snarls x 10,000 haplotypes, binary trait,
chi2 + Fisher, per GPU (weak scaling: every rank tests its own snarl shard, no data-path collective —
SURVEY §8e "snarl-sharded").  A step = one pass of the hot path (K1 path_and_popc + K2 binary_stats) over the
whole resident batch; the 5 GB matrix is far larger than the 126 MB L2, so every step streams it from HBM.

  value     whole-job path tests / s, inputs resident in HBM (matrix repacked, CSR staged)
  e2e       same metric through the host-buffer C-ABI calls (set_matrix + set_binary_phenotype + binary_batch):
            H2D of the reference-layout matrix and the CSR, K0 repack, prep, K1, K2, D2H of every result array
  roofline  K1 (dominant kernel): algorithmic bytes / CUDA-event duration vs the measured HBM copy bandwidth
  cpu_baseline  the oracle (CPU restatement of the reference loop, OpenMP, all host threads) on a bounded sample

--impl reference times that CPU implementation alone (rank 0), same metric / unit / config.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "snarl-path tests/sec (binary chi2+Fisher)"
UNIT = "path-tests/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--snarls", type=int, default=1_000_000, help="snarls per GPU")
    ap.add_argument("--samples", type=int, default=5_000, help="samples (haplotypes = 2x)")
    ap.add_argument("--seed", type=int, default=42, help="generator seed of rank 0 (rank r draws seed + r)")
    ap.add_argument("--e2e-steps", type=int, default=2)
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="target CPU time of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--trait", default="binary", choices=["binary", "quantitative", "eqtl"])
    ap.add_argument("--genes", type=int, default=20_000, help="eqtl: genes (C4: 200k snarls x 5k samples x 20k genes)")
    ap.add_argument("--p-threshold", type=float, default=1e-5, help="eqtl: significance threshold of the scan summary")
    args = ap.parse_args()
    if args.trait == "eqtl" and "--snarls" not in " ".join(sys.argv):
        args.snarls = 200_000
    return args


def config_dict(args, n_gpus):
    return {
        "workload": f"C3 synthetic {args.snarls} snarls x {2 * args.samples} haplotypes {args.trait} "
                    f"({'chi2+Fisher' if args.trait == 'binary' else 'OLS'}), per GPU",
        "snarls_per_gpu": args.snarls, "haplotypes": 2 * args.samples,
        "sharding": "snarl-sharded, no collective" if n_gpus > 1 else "single GPU",
        "l2": "inputs (5 GB matrix) exceed the 126 MB L2: every step streams from HBM, no flush needed",
        "thresholds": {"min_individuals": 3, "min_haplotypes": 5, "maf": 0.05},
    }


# ---------------------------------------------------------------------------------------------------
# clocks
# ---------------------------------------------------------------------------------------------------
class ClockSampler:
    """SM clock + throttle reasons DURING the timed region.  The region lasts tens of milliseconds, so NVML is polled
    from a thread every 2 ms (nvidia-smi -lms cannot sample that fast); nvidia-smi is the fallback."""
    REASONS = (("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20),
               ("sw_power_cap", 0x4))
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.sm, self.mask, self.max_mhz, self.stop_flag, self.thread = index, [], 0, None, False, None
        self.nvml = None
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nvml = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.nvml = None

    def _poll(self):
        n = self.nvml
        get_reasons = getattr(n, "nvmlDeviceGetCurrentClocksEventReasons", None) or \
            getattr(n, "nvmlDeviceGetCurrentClocksThrottleReasons")
        while not self.stop_flag:
            try:
                self.sm.append(float(n.nvmlDeviceGetClockInfo(self.h, n.NVML_CLOCK_SM)))
                self.mask |= int(get_reasons(self.h))
            except Exception:
                pass
            time.sleep(0.002)

    def start(self):
        if self.nvml:
            self.thread = threading.Thread(target=self._poll, daemon=True)
            self.thread.start()

    def stop(self) -> dict:
        if self.nvml:
            self.stop_flag = True
            self.thread.join(timeout=1.0)
            reasons = [name for name, bit in self.REASONS if self.mask & bit]
            return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.max_mhz,
                    "reasons": reasons, "samples": len(self.sm), "source": "nvml, 2 ms poll over the timed region"}
        return self._smi_once()

    def _smi_once(self) -> dict:
        try:
            out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i",
                                  str(self.index)], capture_output=True, text=True, timeout=10).stdout.strip()
            f = [x.strip() for x in out.split(",")]
            reasons = [n for n, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"),
                                         f[5:9]) if v.lower().startswith("active")]
            return {"sm_mhz": float(f[1]), "sm_max_mhz": float(f[2]), "reasons": reasons, "samples": 1,
                    "source": "nvidia-smi, one query right after the timed region (nvml unavailable)"}
        except Exception:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"], "samples": 0}


# ---------------------------------------------------------------------------------------------------
# CPU arm (oracle = port of the reference loop)
# ---------------------------------------------------------------------------------------------------
def cpu_sample_dataset(n_snarls, n_samples, seed=42):
    from stoat_b200 import synth

    return synth.numpy_dataset(n_snarls, n_samples, seed=seed)


def cpu_kind():
    """"reference": oracle/_ref (the reference's own analyzer sources, built by oracle/ref_build.py, shipped to the box as a
    prebuilt .so) when it is there; "port": the oracle restatement otherwise."""
    try:
        from oracle import ref

        if os.path.exists(ref.ref_build.LIB) or ref.available():
            ref.lib()
            return "reference"
    except Exception:
        pass
    return "port"


def cpu_threads(kind):
    import oracle

    if kind == "reference":
        from oracle import ref

        return ref.num_threads()
    return oracle.num_threads()


def cpu_run(d, trait, threads, passes=1, kind="port"):
    """`passes` runs of the CPU snarl loop (the reference's own analyzers, or the oracle port) over the sample;
    -> (seconds of the loop as the reference's "Snarl time analysis" timer spans it, path tests)."""
    import oracle

    secs = 0.0
    tsv = os.path.join("/tmp", f"stoat_cpu_{os.getpid()}.tsv")
    if kind == "reference":
        from oracle import ref

        nodes = d.get("_nodes")
        if nodes is None:
            nodes = d["_nodes"] = ref.nodes_from_structure(d["structure"])
        for _ in range(passes):
            r = ref.run(trait, d["bits"], d["n_rows"], d["n_cols"], d["paths"].snarl_path_off, nodes,
                        is_case=d["is_case"], y=d["y"], n_threads=threads, tsv_path=tsv, parse=False)
            secs += r["seconds"]
        return secs, passes * d["paths"].n_paths
    P = oracle.Paths(d["paths"].snarl_path_off, d["paths"].path_edge_off, d["paths"].edge_row)
    for _ in range(passes):
        if trait == "binary":
            r = oracle.binary_batch(d["bits"], d["n_cols"], d["is_case"], P, n_threads=threads, tsv_path=tsv)
        else:
            r = oracle.quantitative_batch(d["bits"], d["n_cols"], d["y"], P, n_threads=threads)
        secs += r["seconds"]
    return secs, passes * d["paths"].n_paths


# the host-side generator holds the sample as a dense bool matrix: 20,000 snarls x 10,000 haplotypes is
# ~0.8 GB, so longer CPU runs repeat passes over that sample (its 100 MB bit matrix exceeds the host caches)
CPU_SAMPLE_MAX_SNARLS = 20000


def cpu_sample_plan(rate, target_seconds):
    """(snarls in the sample, passes over it) for about `target_seconds` of CPU work at `rate` tests/s."""
    want = max(400, int(rate * target_seconds / 2.05))
    n = min(CPU_SAMPLE_MAX_SNARLS, want)
    passes = max(1, int(round(want / n)))
    return n, passes


def best_threads(probe, trait, kind):
    """All host threads unless the box's CPU quota makes one thread faster (cgroup-limited containers)."""
    best = (1, 0.0)
    for th in sorted({1, cpu_threads(kind)}):
        secs, tests = cpu_run(probe, trait, th, kind=kind)
        if tests / max(secs, 1e-9) > best[1]:
            best = (th, tests / max(secs, 1e-9))
    return best


CPU_WHAT = {"reference": "the reference's own analyzer sources (oracle/_ref: src/snarl_analyzer.cpp loop, tables, "
                         "FisherKhi2 / Eigen OLS, writer; Boost tails from stand-ins) incl. per-row TSV write",
            "port": "oracle port of the reference's OpenMP snarl loop incl. per-row TSV write"}


def cpu_baseline(args, target_seconds):
    """Time the CPU implementation on a bounded sample of the same workload (same generator, same shapes)."""
    kind = cpu_kind()
    probe = cpu_sample_dataset(200, args.samples)
    threads, rate = best_threads(probe, args.trait, kind)
    n, passes = cpu_sample_plan(rate, target_seconds)
    d = cpu_sample_dataset(n, args.samples)
    secs, tests = cpu_run(d, args.trait, threads, passes, kind=kind)
    return {"value": tests / secs, "unit": UNIT, "cores": threads, "kind": kind,
            "sample": f"{passes} passes over the first {n} snarls ({tests} path tests in all) of the same generator, "
                      f"{2 * args.samples} haplotypes, {CPU_WHAT[kind]}, {secs:.2f} s"}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import oracle

    oracle.build()
    kind = cpu_kind()
    probe = cpu_sample_dataset(200, args.samples)
    threads, rate = best_threads(probe, args.trait, kind)
    # each step ~ 4 s of CPU work so that warmup + steps stay within a few minutes
    n, passes = cpu_sample_plan(rate, 4.0)
    d = cpu_sample_dataset(n, args.samples)
    for _ in range(args.warmup):
        cpu_run(d, args.trait, threads, passes, kind=kind)
    total = 0.0
    tests = 0
    for _ in range(args.steps):
        s, t = cpu_run(d, args.trait, threads, passes, kind=kind)
        total += s
        tests += t
    value = tests / total
    sample = (f"{passes} passes over {n} snarls ({passes * d['paths'].n_paths} path tests) per step, same generator "
              f"and shapes, {CPU_WHAT[kind]}")
    line = {
        "impl": "reference", "metric": METRIC if args.trait == "binary" else "snarl-path tests/sec (quantitative OLS)",
        "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u32 popcounts + f64 statistics", "data": "synthetic",
        "config": config_dict(args, args.gpus),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


# ---------------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------------
def algorithmic_bytes_k1(paths, pitch):
    """sum_s U_s * W + W (case mask) + 4 * sum R (edge_row) + 8 * (S + P) (offsets) + 8 * P (g0, g1)."""
    er = paths.edge_row
    S, P = paths.n_snarls, paths.n_paths
    # distinct rows per snarl: rows are private to a snarl in the generator, so distinct over the batch
    distinct = np.unique(er[er != 0xFFFFFFFF]).size
    return distinct * pitch + pitch + 4 * er.size + 8 * (S + P) + 8 * P


def algorithmic_bytes_k3m(paths, pitch):
    """K3M touches the rows of the two-path snarls only (the others go to the left-over list): sum over those snarls of
    U_s * W, + 16 B lean record in, + 52 B partial sums and 1 B state out per snarl."""
    spo = np.asarray(paths.snarl_path_off, dtype=np.int64)
    peo = np.asarray(paths.path_edge_off, dtype=np.int64)
    er = np.asarray(paths.edge_row)
    two = np.nonzero(np.diff(spo) == 2)[0]
    e0, e1 = peo[spo[two]], peo[spo[two + 1]]
    take = np.zeros(er.size + 1, np.int64)
    np.add.at(take, e0, 1)
    np.add.at(take, e1, -1)
    sel = er[np.cumsum(take[:-1]) > 0]
    distinct = np.unique(sel[sel != 0xFFFFFFFF]).size
    return distinct * pitch + (16 + 53) * paths.n_snarls


def bind_to_gpu_numa(local: int) -> None:
    """One process per GPU: run on the CPUs next to that GPU (NVML's ideal affinity) so that the pinned host buffers of the
    end-to-end leg are first-touched on its NUMA node — with 4 or 8 ranks on the default affinity the host-to-device copies
    of all ranks cross the socket interconnect and e2e stops scaling (measured: 111 GB/s aggregate at N = 4)."""
    if os.environ.get("STOAT_BENCH_NO_BIND"):
        return
    try:
        import pynvml

        pynvml.nvmlInit()
        pynvml.nvmlDeviceSetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(local))
    except Exception as e:  # no NVML / not permitted: keep the inherited affinity
        print(f"[bench] cpu affinity not set for GPU {local}: {e}", file=sys.stderr)


def quant_traffic():
    """dram bytes per launch of K3M from the committed ncu capture (profiles/k3m_traffic.json), or None"""
    try:
        with open(os.path.join(ROOT, "profiles", "k3m_traffic.json")) as f:
            return json.load(f).get("dram_bytes_per_launch")
    except Exception:
        return None


def main():
    args = parse_args()
    if args.trait == "eqtl":
        import bench_eqtl

        bench_eqtl.run(args, ClockSampler)
        return
    if args.impl == "reference":
        run_reference_arm(args)
        return

    import torch
    import torch.distributed as dist

    import stoat_b200
    from stoat_b200 import synth

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        bind_to_gpu_numa(local)
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200: stoat_b200 has no CPU fallback (use --impl reference for the CPU arm)")
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- data: drawn on the device, one independent shard per rank (weak scaling) ------------------
    d = synth.torch_dataset(args.snarls, args.samples, dev, seed=args.seed + rank)
    paths = d["paths"]
    S, P = paths.n_snarls, paths.n_paths
    ctx = stoat_b200.Context(local)
    ctx.set_matrix_dev(d["bits_dev"].data_ptr(), d["n_rows"], d["n_cols"])
    t_repack = ctx.timing()["repack_ms"]
    pitch = ((int(d["n_cols"]) + 127) // 128) * 16  # SURVEY §8d W (the device pitch is rounded up to whole 32 B sectors)
    if args.trait == "binary":
        ctx.set_binary_phenotype(d["is_case"])
        run = ctx.binary_run_staged
    else:
        ctx.set_quantitative_phenotype(d["y"])
        run = ctx.quantitative_run_staged
    ctx.stage_paths(paths)
    stream = torch.cuda.ExternalStream(ctx.stream, device=dev)

    for _ in range(max(args.warmup, 3)):
        run()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    k1_ms, k2_ms, launches, n_general, n_fisher = [], [], 0, 0, 0
    e0.record(stream)
    for _ in range(args.steps):
        run()
        t = ctx.timing()
        k1_ms.append(t["table_ms"])
        k2_ms.append(t["stats_ms"])
        launches += t["launches"]
        n_general, n_fisher = t["n_general"], t["n_fisher"]
    e1.record(stream)
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    elapsed_ms = e0.elapsed_time(e1)
    if os.environ.get("STOAT_BENCH_DEBUG"):
        print(f"[rank {rank}] elapsed {elapsed_ms:.3f} ms over {args.steps} steps; kernels {np.sum(k1_ms) + np.sum(k2_ms):.3f} ms; "
              f"per step table {np.round(k1_ms, 3).tolist()} stats {np.round(k2_ms, 3).tolist()}", file=sys.stderr, flush=True)
    t_max = torch.tensor([elapsed_ms], dtype=torch.float64, device=dev)
    launches_t = torch.tensor([launches], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t_max, op=dist.ReduceOp.MAX)
        dist.all_reduce(launches_t, op=dist.ReduceOp.SUM)
    elapsed_ms = float(t_max.item())
    value = world * P * args.steps / (elapsed_ms * 1e-3)

    # ---- end to end: host buffers through the drop-in calls ----------------------------------------
    e2e = None
    if not args.no_e2e:
        host_bits = torch.empty(d["bits_dev"].numel(), dtype=torch.uint8, pin_memory=True)
        host_bits.copy_(d["bits_dev"])
        torch.cuda.synchronize()
        if args.trait == "binary":
            out = ctx._binary_arrays(S, P)
        else:
            out = ctx._quant_arrays(S, P)

        def e2e_step():
            ctx.set_matrix_ptr(host_bits.data_ptr(), d["n_rows"], d["n_cols"])
            if args.trait == "binary":
                ctx.set_binary_phenotype(d["is_case"])
                ctx.binary_batch(paths, out=out)
            else:
                ctx.set_quantitative_phenotype(d["y"])
                ctx.quantitative_batch(paths, out=out)

        e2e_step()  # warm-up (allocations)
        barrier()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record(stream)
        for _ in range(args.e2e_steps):
            e2e_step()
        f1.record(stream)
        barrier()
        te = torch.tensor([f0.elapsed_time(f1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        h2d = host_bits.numel() + paths.snarl_path_off.nbytes + paths.path_edge_off.nbytes + paths.edge_row.nbytes + \
            (d["is_case"].nbytes if args.trait == "binary" else d["y"].nbytes)
        d2h = sum(v.nbytes for v in out.values())
        e2e = {"value": world * P * args.e2e_steps / (float(te.item()) * 1e-3), "unit": UNIT,
               "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
               "ms_per_step": float(te.item()) / args.e2e_steps,
               "includes": "set_matrix (H2D + K0 repack) + phenotype + stage (H2D CSR, prep) + K1 + K2 + D2H results"}
        del host_bits

    if rank == 0:
        peaks = {}
        try:
            with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
                peaks = json.load(f)
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6.65 TB/s (B200_PROFILING.md)"
        k1 = float(np.mean(k1_ms))
        alg = algorithmic_bytes_k1(paths, pitch) if args.trait == "binary" else algorithmic_bytes_k3m(paths, pitch)
        achieved = alg / (k1 * 1e-3) / 1e9
        traffic = None
        try:
            with open(os.path.join(ROOT, "profiles", "k1_traffic.json")) as f:
                traffic = json.load(f).get("dram_bytes_per_launch")
        except Exception:
            pass
        roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                    "traffic": traffic if args.trait == "binary" else quant_traffic(),
                    "kernel": "path_and_popc (K1)" if args.trait == "binary" else "dosage_moments_mma (K3M)", "algorithmic_bytes_per_launch": int(alg),
                    "kernel_ms": k1, "peak_source": peak_src,
                    "snarls_off_fast_path": int(n_general), "fisher_tables": int(n_fisher),
                    "other_kernels_ms": {"binary_stats (K2a) + fisher_exact (K2b)" if args.trait == "binary"
                                         else "three-path / thirds snarls (K3M3, K3M2x) + left-over list (K3) + ols_solve, student_tail (K4a, K4b) + general path (K3G)": float(np.mean(k2_ms)),
                                         "repack_rows (K0, once per matrix, not in the step)": t_repack}}
        line = {
            "metric": METRIC if args.trait == "binary" else "snarl-path tests/sec (quantitative OLS)",
            "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": elapsed_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u32 popcounts + f64 statistics", "data": "synthetic (seed 42, drawn on device)",
            "config": config_dict(args, world), "clocks": clocks, "e2e": e2e,
            "gpu_launches": int(launches_t.item()), "roofline": roofline,
        }
        if world == 1 and not args.no_cpu_baseline:
            import oracle

            oracle.build()
            line["cpu_baseline"] = cpu_baseline(args, args.cpu_seconds)
        print(json.dumps(line))
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
