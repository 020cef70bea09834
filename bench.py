#!/usr/bin/env python3
"""bench.py — snarl-path tests/sec of the `stoat vcf` association hot path on B200 (BASELINE.json's metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
  torchrun --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...        (one rank per GPU)

Headline (no flags): BASELINE.json configs[2] "C3" — synthetic 1,000,000 snarls x 10,000 haplotypes per GPU, binary
trait, chi2 + Fisher.  Weak scaling: every rank tests its own snarl shard, no data-path collective (SURVEY §8e
"snarl-sharded").  A step = one pass of the hot path over the whole resident batch; the 5 GB matrix is far larger than
the 126 MB L2, so every step streams it from HBM.  The same run also measures the other two traits the metric names
and nests them under "traits": quantitative (C3, OLS per snarl) and eqtl (C4: 200k snarls x 5k samples x 20k genes).

  value     whole-job path tests / s, inputs resident in HBM (matrix repacked, CSR staged); the K steps are enqueued
            back to back on the context's stream, CUDA events around them, max over ranks
  e2e       same metric through the host-buffer C-ABI calls (set_matrix + set_*_phenotype + *_batch): H2D of the
            reference-layout matrix and the CSR from pinned host memory, K0 repack, prep, K1, K2, D2H of every result
  roofline  the count pass K1 path_and_popc (binary) / K3M dosage_moments_mma (quantitative): algorithmic bytes / mean
            CUDA-event duration over the timed steps vs the measured HBM copy bandwidth.  `step_frac` is the same bytes
            over the WHOLE step (count pass + statistics), i.e. how far the step as a whole is from the HBM roof;
            `dominant_kernel` names whichever part of the step takes longer
  parity    a spot check ON THE TIMED DATA: the first snarls of the batch, their matrix rows read back from the device,
            through the CPU oracle; mismatches must be 0
  cpu_baseline  the reference's own analyzer sources (oracle/_ref, kind "reference"; the oracle port where that .so is
            absent) on a bounded sample of the same generator, all host threads

  --config c5 [--sharding snarl|sample]   BASELINE.json configs[4] "C5": 10M snarls x 50,000 haplotypes over 8 GPUs =
            1.25M snarls x 50k haplotypes per GPU.  snarl: every rank its own snarl shard (weak scaling, no collective).
            sample: the ranks split the haplotype columns of ONE 1.25M-snarl shard and all-reduce the u32 count buffer
            (NCCL, in place, inside the timed step) between the count pass and the statistics (strong scaling).
  --impl reference   the CPU implementation alone (rank 0), same metric / unit / config.
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

UNIT = "path-tests/s"
METRICS = {"binary": "snarl-path tests/sec (binary chi2+Fisher)", "quantitative": "snarl-path tests/sec (quantitative OLS)"}
CONFIGS = {"c3": dict(snarls=1_000_000, samples=5_000), "c5": dict(snarls=1_250_000, samples=25_000)}


def parse_args(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="c3", choices=sorted(CONFIGS))
    ap.add_argument("--sharding", default="snarl", choices=["snarl", "sample", "gene"])
    ap.add_argument("--snarls", type=int, default=0, help="snarls per GPU (default: the config's)")
    ap.add_argument("--samples", type=int, default=0, help="samples (haplotypes = 2x; default: the config's)")
    ap.add_argument("--seed", type=int, default=42, help="generator seed of rank 0 (rank r draws seed + r)")
    ap.add_argument("--e2e-steps", type=int, default=2)
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="target CPU time of a cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--parity-snarls", type=int, default=20_000)
    ap.add_argument("--trait", default="binary", choices=["binary", "quantitative", "eqtl"])
    ap.add_argument("--traits", default="all", choices=["all", "headline"],
                    help="all: the C3 binary headline also carries the quantitative and eQTL records")
    ap.add_argument("--genes", type=int, default=20_000, help="eqtl: genes (C4: 200k snarls x 5k samples x 20k genes)")
    ap.add_argument("--ncov", type=int, default=0,
                    help="quantitative / eqtl: covariates (SURVEY §8d 'then ncov = 3'; every snarl then takes the general kernel)")
    ap.add_argument("--p-threshold", type=float, default=1e-5, help="eqtl: significance threshold of the scan summary")
    args = ap.parse_args(argv)
    args.custom = bool(args.snarls or args.samples)
    if not args.snarls:
        args.snarls = 200_000 if args.trait == "eqtl" else CONFIGS[args.config]["snarls"]
    if not args.samples:
        args.samples = 5_000 if args.trait == "eqtl" else CONFIGS[args.config]["samples"]
    return args


def config_dict(args, n_gpus, trait):
    name = "custom" if args.custom else args.config.upper()
    what = "chi2+Fisher" if trait == "binary" else "OLS"
    mat_gb = args.snarls * 4.07 * (2 * args.samples / 8) / 1e9
    if args.sharding == "sample" and n_gpus > 1:
        shard = (f"sample-sharded: {n_gpus} ranks split the {2 * args.samples} haplotype columns of one {args.snarls}-snarl "
                 f"shard; ncclAllReduce(sum, u32, in place) of the g0|g1 count buffer between K1 and K2 inside every step")
    elif n_gpus > 1:
        shard = "snarl-sharded: every rank its own snarl shard, no collective"
    else:
        shard = "single GPU"
    return {
        "workload": f"{name} synthetic {args.snarls} snarls x {2 * args.samples} haplotypes {trait} ({what}), per GPU",
        "snarls_per_gpu": args.snarls, "haplotypes": 2 * args.samples, "sharding": shard,
        "l2": f"inputs ({mat_gb:.1f} GB matrix) exceed the 126 MB L2: every step streams from HBM, no flush needed",
        "thresholds": {"min_individuals": 3, "min_haplotypes": 5, "maf": 0.05},
        "covariates": (f"{args.ncov} covariates: every snarl on the general kernel (K3G); the integer-MMA kernel the roofline "
                       "object names does not run" if (getattr(args, "ncov", 0) and trait == "quantitative") else "none"),
    }


def bench_covariates(args, n_samples):
    """Seeded covariate matrix of a --ncov run [n_samples][ncov] (None without)."""
    if not getattr(args, "ncov", 0):
        return None
    import numpy as np

    return np.random.default_rng(11).normal(size=(n_samples, args.ncov))


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
# CPU arm: the reference's own analyzer loop (oracle/_ref), or the oracle port where that .so is absent
# ---------------------------------------------------------------------------------------------------
def host_threads() -> int:
    """Host threads this process may use: its CPU affinity, capped by the cgroup quota.  NOT omp_get_max_threads():
    torchrun exports OMP_NUM_THREADS=1, which made the round-1 reference arm run on one core at N >= 2."""
    try:
        n = len(os.sched_getaffinity(0))
    except AttributeError:
        n = os.cpu_count() or 1
    try:
        with open("/sys/fs/cgroup/cpu.max") as f:
            quota, period = f.read().split()
        if quota != "max":
            n = max(1, min(n, int(float(quota) / float(period) + 0.5)))
    except Exception:
        pass
    return max(1, n)


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


def cpu_sample_plan(rate, target_seconds, n_samples):
    """(snarls in the sample, passes over it) for about `target_seconds` of CPU work at `rate` tests/s.  The host-side
    generator holds the sample as a dense bool matrix (4 rows per snarl x haplotypes): capped at ~0.8 GB, longer runs repeat
    passes over the sample (its bit matrix, >= 100 MB, exceeds the host caches)."""
    cap = max(400, int(8e8 / (4.07 * 2 * n_samples)))
    want = max(400, int(rate * target_seconds / 2.05))
    n = min(cap, want)
    passes = max(1, int(round(want / n)))
    return n, passes


def best_threads(probe, trait, kind):
    """All host threads unless the box's CPU quota makes one thread faster (cgroup-limited containers)."""
    best = (1, 0.0)
    for th in sorted({1, host_threads()}):
        secs, tests = cpu_run(probe, trait, th, kind=kind)
        if tests / max(secs, 1e-9) > best[1]:
            best = (th, tests / max(secs, 1e-9))
    return best


CPU_WHAT = {"reference": "the reference's own analyzer sources (oracle/_ref: src/snarl_analyzer.cpp loop, tables, "
                         "FisherKhi2 / Eigen OLS, writer; Boost tails from stand-ins) incl. per-row TSV write",
            "port": "oracle port of the reference's OpenMP snarl loop incl. per-row TSV write"}


def cpu_baseline(args, trait, target_seconds):
    """Time the CPU implementation on a bounded sample of the same workload (same generator, same shapes)."""
    kind = cpu_kind()
    probe = cpu_sample_dataset(200, args.samples)
    threads, rate = best_threads(probe, trait, kind)
    n, passes = cpu_sample_plan(rate, target_seconds, args.samples)
    d = cpu_sample_dataset(n, args.samples)
    secs, tests = cpu_run(d, trait, threads, passes, kind=kind)
    return {"value": tests / secs, "unit": UNIT, "cores": threads, "kind": kind,
            "sample": f"{passes} passes over the first {n} snarls ({tests} path tests in all) of the same generator, "
                      f"{2 * args.samples} haplotypes, {CPU_WHAT[kind]}, {secs:.2f} s"}


def reference_record(args, trait, steps, warmup):
    """The CPU implementation of one presence-matrix trait, `steps` bounded samples of ~4 s each -> its JSON record."""
    kind = cpu_kind()
    probe = cpu_sample_dataset(200, args.samples)
    threads, rate = best_threads(probe, trait, kind)
    # each step ~ 4 s of CPU work so that warmup + steps stay within a few minutes
    n, passes = cpu_sample_plan(rate, 4.0, args.samples)
    d = cpu_sample_dataset(n, args.samples)
    for _ in range(warmup):
        cpu_run(d, trait, threads, passes, kind=kind)
    total = 0.0
    tests = 0
    for _ in range(steps):
        s, t = cpu_run(d, trait, threads, passes, kind=kind)
        total += s
        tests += t
    value = tests / total
    sample = (f"{passes} passes over {n} snarls ({passes * d['paths'].n_paths} path tests) per step, same generator "
              f"and shapes, {CPU_WHAT[kind]}")
    return {
        "impl": "reference", "metric": METRICS[trait], "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": steps, "warmup": warmup, "ms_per_step": 1e3 * total / steps, "higher_is_better": True,
        "scaling": "strong" if args.sharding == "sample" else "weak",
        "vs_baseline": None, "dtype": "u32 popcounts + f64 statistics", "data": "synthetic",
        "config": config_dict(args, args.gpus, trait),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }


def run_reference_arm(args):
    """Rank 0 alone times the reference's CPU implementation (all the host threads it may use); other ranks exit."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import oracle

    oracle.build()
    if args.trait == "eqtl":
        import bench_eqtl

        bench_eqtl.run_reference_arm(args, host_threads())
        return
    line = reference_record(args, args.trait, args.steps, args.warmup)
    if args.traits == "all" and args.trait == "binary" and args.config == "c3" and not args.custom:
        import bench_eqtl

        eargs = parse_args(["--trait", "eqtl", "--gpus", str(args.gpus), "--steps", "2", "--warmup", "1"])
        line["traits"] = {"quantitative": reference_record(args, "quantitative", 2, 1),
                          "eqtl": bench_eqtl.reference_line(eargs, host_threads())}
    print(json.dumps(line))


# ---------------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------------
def survey_pitch(n_cols):
    """SURVEY §8d's W = ceil(H / 128) * 16: the unit of the algorithmic bytes (the device pitch rounds up to 32-byte sectors)."""
    return ((int(n_cols) + 127) // 128) * 16


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


def committed_traffic(name):
    """dram bytes per launch from the committed ncu capture (profiles/<name>), or None"""
    try:
        with open(os.path.join(ROOT, "profiles", name)) as f:
            return json.load(f).get("dram_bytes_per_launch")
    except Exception:
        return None


def load_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
        if "hbm_gbs" in peaks:
            return float(peaks["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        pass
    return 6650.0, "fallback 6.65 TB/s (B200_PROFILING.md)"


def mem_available_bytes():
    try:
        with open("/proc/meminfo") as f:
            for ln in f:
                if ln.startswith("MemAvailable:"):
                    return int(ln.split()[1]) * 1024
    except Exception:
        pass
    return None


class Env:
    """torch / torch.distributed state of this rank"""

    def __init__(self):
        import torch
        import torch.distributed as dist

        self.torch, self.dist = torch, dist
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py needs a B200: stoat_b200 has no CPU fallback (use --impl reference for the CPU arm)")
        if self.world > 1:
            bind_to_gpu_numa(self.local)
        self.dev = torch.device("cuda", self.local)
        torch.cuda.set_device(self.dev)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=self.dev)

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, v: float) -> float:
        t = self.torch.tensor([v], dtype=self.torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(self, v: float) -> float:
        t = self.torch.tensor([v], dtype=self.torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM)
        return float(t.item())

    def close(self):
        if self.world > 1:
            self.dist.destroy_process_group()


def parity_spot_check(env, ctx, d, trait, n_check, res, col_sharded, covar=None):
    """The first `n_check` snarls of the TIMED batch through the CPU oracle: their matrix rows come back from the device
    copy the kernels read (reference layout, byte-aligned rows), the results are the ones the timed configuration produced.
    Sample-sharded runs gather every rank's column slice of those rows first.  -> dict for the JSON line (rank 0)."""
    import oracle

    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import helpers

    torch = env.torch
    st = d["structure"]
    S = d["paths"].n_snarls
    n = min(n_check, S)
    r1 = int(st.row0[n - 1] + st.n_rows_of[n - 1])
    row_bytes = d["n_cols"] // 8
    pre = d["bits_dev"][: r1 * row_bytes].view(r1, row_bytes)
    is_case, y = d["is_case"], d["y"]
    n_cols = d["n_cols"]
    if col_sharded and env.world > 1:
        # the ranks' column slices differ in width (the last one is what is left): pad to the widest, gather, trim
        widths = [torch.zeros(1, dtype=torch.int64, device=pre.device) for _ in range(env.world)]
        env.dist.all_gather(widths, torch.tensor([row_bytes], dtype=torch.int64, device=pre.device))
        widths = [int(w.item()) for w in widths]
        wmax = max(widths)
        mine = torch.zeros((r1, wmax), dtype=pre.dtype, device=pre.device)
        mine[:, :row_bytes] = pre
        parts = [torch.empty_like(mine) for _ in range(env.world)]
        env.dist.all_gather(parts, mine)
        pre = torch.cat([p_[:, :w] for p_, w in zip(parts, widths)], dim=1)
        n_cols = 8 * sum(widths)
        is_case, y = d["is_case_all"], d["y_all"]
    if env.rank != 0:
        return None
    bits = pre.contiguous().cpu().numpy().reshape(-1)
    p = d["paths"]
    P1, E1 = int(p.snarl_path_off[n]), int(p.path_edge_off[int(p.snarl_path_off[n])])
    sub = oracle.Paths(p.snarl_path_off[: n + 1].copy(), p.path_edge_off[: P1 + 1].copy(), p.edge_row[:E1].copy())
    t0 = time.time()
    if trait == "binary":
        ref = oracle.binary_batch(bits, n_cols, is_case, sub, n_threads=host_threads())
        got = {k: (v[:P1] if k in ("g0", "g1") else v[:n]) for k, v in res.items()}
        check = helpers.assert_binary_equal
    else:
        ref = oracle.quantitative_batch(bits, n_cols, y, sub, covar=covar, n_threads=host_threads())
        got = {k: (v[:P1] if k == "allele_paths" else v[:n]) for k, v in res.items()}
        check = helpers.assert_quant_equal
    try:
        check(got, ref)
        mism, note = 0, None
    except AssertionError as e:
        mism, note = 1, str(e)[:300]
    out = {"checked_snarls": int(n), "checked_path_tests": int(P1), "kept_rows": int((ref["filtered"] == 0).sum()),
           "mismatches": mism, "against": "oracle port (oracle/stoat_oracle.cpp) on the timed matrix's own rows; tables / "
           "flags bit-exact, p-values / betas 1e-9 relative", "seconds": round(time.time() - t0, 2)}
    if note:
        out["first_mismatch"] = note
    return out


def logit_record(env, args, ctx, d, stream):
    """The binary trait WITH covariates (BinaryCovarSnarlAnalyzer: a logistic regression per snarl) on the
    resident batch: device-resident timing of stoat_gpu_binary_covar_run_staged, parity of the first snarls against the
    oracle.  As the reference does, the covariates have no effect on the fit (src/stats_test.cpp:59-62)."""
    torch = env.torch
    paths = d["paths"]
    ctx.set_binary_phenotype(d["is_case"])
    ctx.stage_paths(paths)
    steps, warm = args.steps, max(args.warmup, 3)
    for _ in range(warm):
        ctx.binary_covar_run_staged()
    env.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(steps):
        ctx.binary_covar_run_staged()
    e1.record(stream)
    env.barrier()
    elapsed_ms = env.max_over_ranks(e0.elapsed_time(e1))
    res = ctx.binary_covar_fetch()
    t = ctx.timing()
    launches = int(env.sum_over_ranks(t["launches"] * steps))
    if env.rank != 0:
        return None
    parity = None
    if not args.no_parity:
        import oracle

        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import test_gpu_logit

        k = min(paths.n_snarls, 5000)
        st = d["structure"]
        r1 = int(st.row0[k - 1] + st.n_rows_of[k - 1])
        bits = d["bits_dev"][: r1 * (d["n_cols"] // 8)].cpu().numpy()
        P1 = int(paths.snarl_path_off[k])
        E1 = int(paths.path_edge_off[P1])
        sub = oracle.Paths(paths.snarl_path_off[: k + 1].copy(), paths.path_edge_off[: P1 + 1].copy(), paths.edge_row[:E1].copy())
        ref = oracle.binary_covar_batch(bits, d["n_cols"], d["is_case"], sub, n_threads=host_threads())
        got = {kk: (v[:P1] if kk == "allele_paths" else v[:k]) for kk, v in res.items()}
        try:
            parity = {"checked_snarls": int(k), "fits": int(test_gpu_logit.assert_logit_equal(got, ref)), "mismatches": 0,
                      "against": "oracle port: tables / flags bit-exact, beta / se / p 1e-9 relative"}
        except AssertionError as e:
            parity = {"checked_snarls": int(k), "mismatches": 1, "first_mismatch": str(e)[:300]}
    return {"metric": "snarl-path tests/sec (binary with covariates: logistic regression)", "unit": UNIT,
            "value": env.world * paths.n_paths * steps / (elapsed_ms * 1e-3), "ms_per_step": elapsed_ms / steps, "steps": steps,
            "n_gpus": env.world, "dtype": "u32 popcounts + f64 IRLS", "gpu_launches": launches,
            "kernels_ms": {"class counts (L1)": t["table_ms"], "IRLS (L2), three-path snarls (L3), general kernel": t["stats_ms"]},
            "snarls_on_general_kernel": int(t["n_general"]), "parity": parity}


def trait_record(env, args, ctx, d, trait, stream, host_cache, with_clocks=True):
    """Measure one presence-matrix trait (binary | quantitative) on the resident data set -> the record of the JSON line."""
    torch = env.torch
    world, rank = env.world, env.rank
    paths = d["paths"]
    S, P = paths.n_snarls, paths.n_paths
    sample_sharded = args.sharding == "sample" and world > 1
    if trait == "binary":
        ctx.set_binary_phenotype(d["is_case"])
    elif sample_sharded:
        from stoat_b200 import shard as shard_mod

        # the local slice + the WHOLE phenotype's mean / centred sums / max |y - mean| (the centring and the fixed-point scale of
        # the integer contraction must be global)
        ctx.set_quantitative_phenotype_shard(d["y"], *shard_mod.phenotype_summary(d["y_all"]))
    else:
        ctx.set_quantitative_phenotype(d["y"])
    covar = bench_covariates(args, len(d["y"])) if (trait == "quantitative" and not sample_sharded) else None
    ctx.set_covariates(covar)
    ctx.stage_paths(paths)
    steps, warm = args.steps, max(args.warmup, 3)
    ar_ms, k1_ms_own, k2_ms_own = [], [], []
    ar_bytes = 0
    if sample_sharded:
        from stoat_b200 import shard as shard_mod

        evs = []
        if trait == "binary":
            ctx.binary_counts_staged()
            ptr, n_words = ctx.binary_counts_dev()  # the buffer only grows: the pointer stays valid for every step below
            counts_t = torch.as_tensor(shard_mod._DevU32(ptr, n_words), device=env.dev)  # int32 view (same sums)
            ar_bytes = n_words * 4
            ar_what = "ncclAllReduce(sum, u32, in place) of g0|g1"

            def run():
                with torch.cuda.stream(stream):
                    a, b, c, e = (torch.cuda.Event(enable_timing=True) for _ in range(4))
                    a.record(stream)
                    ctx.binary_counts_staged()
                    b.record(stream)
                    env.dist.all_reduce(counts_t, op=env.dist.ReduceOp.SUM)
                    c.record(stream)
                    ctx.binary_stats_staged()
                    e.record(stream)
                    evs.append((a, b, c, e))
        else:
            # moments of the local columns (K3M: 7 class counts + 3 sums per snarl and haplotype slice) | ONE exchange: the two
            # partial buffers all-reduced in place | filters + OLS on the sums (K4, replicated).  Snarls that need more than the
            # two-path sums come back deferred; they are finished after the timed loop (shard.replay_deferred_on_device).
            ctx.quantitative_moments_staged()
            pi, ni, pd, nd = ctx.quantitative_partials_dev()
            part_i = torch.as_tensor(shard_mod._DevU32(pi, ni), device=env.dev)
            part_d = torch.as_tensor(shard_mod._DevF64(pd, nd), device=env.dev)
            ar_bytes = ni * 4 + nd * 8
            ar_what = "ncclAllReduce(sum, in place) of K3M's partials: u32 class counts [slices][7][S] and f64 sums [slices][3][S]"

            def run():
                with torch.cuda.stream(stream):
                    a, b, c, e = (torch.cuda.Event(enable_timing=True) for _ in range(4))
                    a.record(stream)
                    ctx.quantitative_moments_staged()
                    b.record(stream)
                    env.dist.all_reduce(part_i, op=env.dist.ReduceOp.SUM)
                    env.dist.all_reduce(part_d, op=env.dist.ReduceOp.SUM)
                    c.record(stream)
                    ctx.quantitative_solve_staged()
                    e.record(stream)
                    evs.append((a, b, c, e))
    else:
        run = ctx.binary_run_staged if trait == "binary" else ctx.quantitative_run_staged

    for _ in range(warm):
        run()
    env.barrier()
    if sample_sharded:
        evs.clear()
    sampler = ClockSampler(env.local)
    if rank == 0 and with_clocks:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(steps):
        run()  # asynchronous: the K steps queue up on the context's stream
    e1.record(stream)
    env.barrier()
    clocks = sampler.stop() if (rank == 0 and with_clocks) else None
    elapsed_ms = env.max_over_ranks(e0.elapsed_time(e1))
    if sample_sharded:
        for a, b, c, e in evs:
            k1_ms_own.append(a.elapsed_time(b))
            ar_ms.append(b.elapsed_time(c))
            k2_ms_own.append(c.elapsed_time(e))
        k1, k2 = float(np.mean(k1_ms_own)), float(np.mean(k2_ms_own))
        launches = (3 if trait == "binary" else 4) * steps  # K1, K2a, K2b | lean records, K3M, K4a, K4b (NCCL's kernels not counted)
        t_last = {"n_general": 0, "n_fisher": 0}
        value = P * steps / (elapsed_ms * 1e-3)  # the ranks share ONE batch
    else:
        tm = ctx.mean_timing(steps)
        k1, k2 = tm["table_ms"], tm["stats_ms"]
        t_last = ctx.timing()
        launches = t_last["launches"] * steps
        value = world * P * steps / (elapsed_ms * 1e-3)
    launches = int(env.sum_over_ranks(launches))
    res = ctx.binary_fetch() if trait == "binary" else ctx.quantitative_fetch()  # results of the last timed step
    deferred = None
    if sample_sharded and trait == "quantitative":
        # outside the timed loop: the deferred snarls (three paths, thirds, ...) from the ranks' gathered row slices, on the device
        import stoat_b200
        from stoat_b200 import shard as shard_mod

        todo = shard_mod.deferred_snarls(res)
        deferred = int(todo.size)
        if todo.size:
            H = 2 * args.samples
            per = ((H + world - 1) // world + 127) // 128 * 128  # the column ranges main() dealt out
            ranges = [(min(H, r * per), min(H, (r + 1) * per)) for r in range(world)]

            def allgather_dev(mine):
                w = torch.tensor([mine.shape[1]], dtype=torch.int64, device=mine.device)
                ws = [torch.zeros_like(w) for _ in range(world)]
                env.dist.all_gather(ws, w)
                wmax = max(int(v.item()) for v in ws)
                padded = torch.zeros((mine.shape[0], wmax), dtype=torch.uint8, device=mine.device)
                padded[:, :mine.shape[1]] = mine
                out = [torch.empty_like(padded) for _ in range(world)]
                env.dist.all_gather(out, padded)
                return out

            t0 = time.time()
            rctx = stoat_b200.Context(env.local)
            try:
                replay = shard_mod.replay_deferred_on_device([ctx], ranges, paths, todo, d["y_all"], (3, 5, 0.05), rctx, allgather_dev)
            finally:
                rctx.close()
            res = shard_mod.merge_deferred(res, todo, replay, paths)
            deferred = {"snarls": deferred, "replay_seconds": round(time.time() - t0, 2)}

    # ---- parity spot check on the timed data ---------------------------------------------------------
    parity = None
    if not args.no_parity:
        parity = parity_spot_check(env, ctx, d, trait, args.parity_snarls, res, sample_sharded, covar=covar)

    # ---- end to end: host buffers through the drop-in calls ----------------------------------------
    e2e = None
    if not args.no_e2e and not sample_sharded:
        nbytes = d["bits_dev"].numel()
        avail = mem_available_bytes()
        if avail is not None and avail < 1.5 * nbytes * world + (8 << 30):
            e2e = {"value": None, "unit": UNIT, "skipped": f"host memory: {avail >> 30} GiB available, the pinned copies of "
                   f"{world} x {nbytes >> 30} GiB matrices do not fit safely"}
        else:
            host_bits = host_cache.get("bits")
            if host_bits is None:
                host_bits = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
                host_bits.copy_(d["bits_dev"])
                torch.cuda.synchronize()
                host_cache["bits"] = host_bits
            # result arrays in pinned host memory too (a pageable destination makes the 43 MB D2H a 14 ms staged copy)
            out = ctx._binary_arrays(S, P) if trait == "binary" else ctx._quant_arrays(S, P)
            pinned = {k: torch.from_numpy(v).pin_memory() for k, v in out.items()}
            out = {k: v.numpy() for k, v in pinned.items()}

            def e2e_step():
                ctx.set_matrix_ptr(host_bits.data_ptr(), d["n_rows"], d["n_cols"])
                if trait == "binary":
                    ctx.set_binary_phenotype(d["is_case"])
                    ctx.binary_batch(paths, out=out)
                else:
                    ctx.set_quantitative_phenotype(d["y"])
                    ctx.quantitative_batch(paths, out=out)

            e2e_step()  # warm-up (allocations)
            t_up = ctx.timing()
            env.barrier()
            f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            f0.record(stream)
            for _ in range(args.e2e_steps):
                e2e_step()
            f1.record(stream)
            env.barrier()
            te = env.max_over_ranks(f0.elapsed_time(f1))
            h2d = host_bits.numel() + paths.snarl_path_off.nbytes + paths.path_edge_off.nbytes + paths.edge_row.nbytes + \
                (d["is_case"].nbytes if trait == "binary" else d["y"].nbytes)
            d2h = sum(v.nbytes for v in out.values())
            e2e = {"value": world * P * args.e2e_steps / (te * 1e-3), "unit": UNIT,
                   "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                   "ms_per_step": te / args.e2e_steps,
                   "includes": "set_matrix (H2D in 256 MB chunks overlapping the K0 repack) + phenotype + stage (H2D CSR, "
                               "prep) + K1 + K2 + D2H results", "last_call_timing_ms": {k: round(float(v), 3) for k, v in t_up.items()
                                                                                       if k.endswith("_ms")}}
            # the same H2D with nothing else going on: what the link gives this rank while all ranks copy at once
            stage = torch.empty(min(nbytes, 1 << 30), dtype=torch.uint8, device=env.dev)
            env.barrier()
            g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            g0.record()
            for _ in range(4):
                stage.copy_(host_bits[: stage.numel()], non_blocking=True)
            g1.record()
            env.barrier()
            e2e["bare_h2d_gbs_per_rank"] = 4 * stage.numel() / (env.max_over_ranks(g0.elapsed_time(g1)) * 1e-3) / 1e9
            e2e["bare_h2d_note"] = "4 x cudaMemcpyAsync of 1 GiB from the same pinned buffer, all ranks at once, slowest rank"
            del stage

    if rank != 0:
        return None
    peak, peak_src = load_peak()
    pitch = survey_pitch(d["n_cols"])
    alg = algorithmic_bytes_k1(paths, pitch) if trait == "binary" else algorithmic_bytes_k3m(paths, pitch)
    achieved = alg / (k1 * 1e-3) / 1e9
    step_ms = elapsed_ms / steps
    names = {"binary": ("path_and_popc (K1)", "binary_stats (K2a) + fisher_exact (K2b)"),
             "quantitative": ("dosage_moments_mma (K3M)", "three-path / thirds snarls (K3M3, K3M2x) + left-over list (K3) + "
                              "ols_solve, student_tail (K4a, K4b) + general path (K3G)")}[trait]
    overlapped = trait == "binary" and not sample_sharded and k1 + k2 > 1.15 * step_ms
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": committed_traffic("k1_traffic.json" if trait == "binary" else "k3m_traffic.json"),
                "kernel": names[0], "algorithmic_bytes_per_launch": int(alg), "kernel_ms": k1, "peak_source": peak_src,
                "step_frac": alg / (step_ms * 1e-3) / 1e9 / peak,
                "step_frac_note": "the count pass's algorithmic bytes over the WHOLE step (ms_per_step) vs the same peak",
                "dominant_kernel": names[0] if k1 >= k2 else names[1],
                "snarls_off_fast_path": int(t_last["n_general"]), "fisher_tables": int(t_last["n_fisher"]),
                "other_kernels_ms": {names[1]: k2, "repack_rows (K0, once per matrix, not in the step)": d.get("repack_ms")},
                "overlap": ("the statistics of snarl chunk c run beside the count pass of chunk c + 1 on a second stream: "
                            "kernel_ms + other_kernels_ms exceeds ms_per_step") if overlapped else None}
    rec = {"metric": METRICS[trait], "value": value, "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": warm,
           "ms_per_step": step_ms, "higher_is_better": True, "scaling": "strong" if sample_sharded else "weak",
           "vs_baseline": None, "dtype": "u32 popcounts + f64 statistics",
           "data": f"synthetic (seed {args.seed}, drawn on device)", "config": config_dict(args, world, trait),
           "clocks": clocks, "e2e": e2e, "gpu_launches": launches, "roofline": roofline, "parity": parity}
    if sample_sharded:
        ar = float(np.mean(ar_ms))
        rec["collective"] = {"op": ar_what, "bytes": int(ar_bytes), "ms": ar,
                             "share_of_step": ar / step_ms, "busbw_gbs": 2 * (world - 1) / world * ar_bytes / (ar * 1e-3) / 1e9,
                             "k1_ms_local_columns": k1, "k2_ms_replicated": k2}
        if deferred is not None:
            rec["collective"]["deferred"] = deferred
    if world == 1 and not args.no_cpu_baseline:
        import oracle

        oracle.build()
        rec["cpu_baseline"] = cpu_baseline(args, trait, args.cpu_seconds)
    return rec


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
        return
    env = Env()
    torch = env.torch
    if args.trait == "eqtl":
        import bench_eqtl

        line = bench_eqtl.measure(args, env, ClockSampler, host_threads())
        if env.rank == 0:
            print(json.dumps(line))
        env.close()
        return

    import stoat_b200
    from stoat_b200 import synth

    world, rank = env.world, env.rank
    sample_sharded = args.sharding == "sample" and world > 1
    # ---- data: drawn on the device.  snarl-sharded: one independent shard per rank (weak scaling); sample-sharded: the
    # ranks hold the column slices (multiples of 128 haplotypes) of ONE shard --------------------------------------------
    if sample_sharded:
        H = 2 * args.samples
        per = ((H + world - 1) // world + 127) // 128 * 128
        c0, c1 = min(H, rank * per), min(H, (rank + 1) * per)
        if c1 <= c0:
            raise SystemExit("bench.py --sharding sample: fewer than 128 haplotypes per rank")
        d = synth.torch_dataset(args.snarls, args.samples, env.dev, seed=args.seed, cols=(c0, c1))
        is_case_all, y_all = synth.phenotypes(args.samples, args.seed)
        d["is_case_all"], d["y_all"] = is_case_all, y_all
    else:
        d = synth.torch_dataset(args.snarls, args.samples, env.dev, seed=args.seed + rank)
    ctx = stoat_b200.Context(env.local)
    ctx.set_matrix_dev(d["bits_dev"].data_ptr(), d["n_rows"], d["n_cols"])
    d["repack_ms"] = ctx.timing()["repack_ms"]
    stream = torch.cuda.ExternalStream(ctx.stream, device=env.dev)
    host_cache = {}
    line = trait_record(env, args, ctx, d, args.trait, stream, host_cache)
    nested = args.traits == "all" and args.trait == "binary" and not args.custom and not sample_sharded
    nested_eqtl = nested and args.config == "c3"
    traits = {}
    if nested:
        q = trait_record(env, args, ctx, d, "quantitative", stream, host_cache, with_clocks=False)
        if rank == 0:
            traits["quantitative"] = q
        if nested_eqtl and not args.ncov:
            # SURVEY §8d "then ncov = 3": the same batch with 3 covariates (sweep kernels K3C / K4C); device-resident
            # timing and parity only, the end-to-end and CPU legs are those of the record above
            import copy

            cargs = copy.copy(args)
            cargs.ncov, cargs.no_e2e, cargs.no_cpu_baseline = 3, True, True
            qc = trait_record(env, cargs, ctx, d, "quantitative", stream, host_cache, with_clocks=False)
            if rank == 0:
                qc["roofline"] = None  # (the record's roofline object describes K3M, which does not run with covariates)
                traits["quantitative_3_covariates"] = qc
            lr = logit_record(env, args, ctx, d, stream)
            if rank == 0:
                traits["binary_with_covariates_logistic"] = lr
    host_cache.clear()
    ctx.close()
    del d
    torch.cuda.empty_cache()
    if nested_eqtl:
        import bench_eqtl

        eargs = parse_args(["--trait", "eqtl", "--gpus", str(args.gpus), "--steps", str(min(args.steps, 3)), "--warmup", "3",
                            "--e2e-steps", "1", "--cpu-seconds", str(args.cpu_seconds)] +
                           (["--no-cpu-baseline"] if args.no_cpu_baseline else []) + (["--no-e2e"] if args.no_e2e else []) +
                           (["--no-parity"] if args.no_parity else []))
        e = bench_eqtl.measure(eargs, env, ClockSampler, host_threads(), with_clocks=False)
        if rank == 0:
            traits["eqtl"] = e
        if not args.ncov:
            # ... and the eQTL scan with 3 covariates (E3C), on 2,000 snarls x 2,000 genes of the same generators
            cargs = parse_args(["--trait", "eqtl", "--gpus", str(args.gpus), "--steps", "3", "--warmup", "3", "--ncov", "3",
                                "--snarls", "2000", "--genes", "2000", "--no-cpu-baseline", "--no-e2e"] +
                               (["--no-parity"] if args.no_parity else []))
            ec = bench_eqtl.measure(cargs, env, ClockSampler, host_threads(), with_clocks=False)
            if rank == 0:
                ec["roofline"] = None  # (describes the tensor-core kernel, which does not run with covariates)
                traits["eqtl_3_covariates"] = ec
    if rank == 0:
        if traits:
            line["traits"] = traits
        print(json.dumps(line))
    env.close()


if __name__ == "__main__":
    main()
