#!/usr/bin/env python3
"""Turn the ncu artefacts a gpurun call brought back (gpurun_out/) into the small tracked summaries under profiles/.

  python scripts/summarise_profiles.py r1a          # tag = round + letter of the GPU call

Writes profiles/<tag>_launches_<trait>.csv (kernel, block, grid, ns per launch), profiles/<tag>_<kernel>.txt (selected
raw metrics + opcode mix + top stall reasons of one `ncu --set full` capture) and profiles/k1_traffic.json
(dram bytes per launch of the dominant kernel, read by bench.py for roofline.traffic).
"""
import csv
import json
import os
import subprocess
import sys
from collections import Counter

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "gpurun_out")
PROF = os.path.join(ROOT, "profiles")

KEEP = ("gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "launch__registers_per_thread", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
        "launch__block_size", "launch__grid_size", "launch__shared_mem_per_block_dynamic",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "sm__cycles_elapsed.max",
        "smsp__thread_inst_executed_per_inst_executed.ratio")


def ncu_csv(rep, page, extra=()):
    out = subprocess.run(["ncu", "-i", rep, "--page", page, "--csv", *extra], capture_output=True, text=True).stdout
    return list(csv.reader(out.splitlines()))


def launches(tag, name):
    src = os.path.join(OUT, f"launches_{name}.csv")
    if not os.path.exists(src):
        return
    rows = [r for r in csv.reader(l for l in open(src) if l.startswith('"'))]
    h = rows[0]
    ik, ib, ig, iv = h.index("Kernel Name"), h.index("Block Size"), h.index("Grid Size"), h.index("Metric Value")
    with open(os.path.join(PROF, f"{tag}_launches_{name}.csv"), "w") as f:
        f.write("kernel,block,grid,gpu__time_duration_ns\n")
        tot = Counter()
        for r in rows[1:]:
            k = r[ik].split("(")[0]
            f.write(f'{k},"{r[ib]}","{r[ig]}",{r[iv]}\n')
            tot[k] += int(float(r[iv].replace(",", "")))
        f.write("# share of the listed launches (ncu serialises and runs cold-cache: compare shares, not absolutes)\n")
        s = sum(tot.values())
        for k, v in tot.most_common():
            f.write(f"# {k},{v} ns,{100.0 * v / s:.1f}%\n")


def full(tag, name):
    rep = os.path.join(OUT, f"prof_{name}.ncu-rep")
    if not os.path.exists(rep):
        return None
    raw = ncu_csv(rep, "raw")
    h, u, v = raw[0], raw[1], raw[2]
    lines = [f"# ncu --set full --clock-control none, one launch of {v[h.index('Kernel Name')]}", ""]
    vals = {}
    for k in KEEP:
        if k in h:
            i = h.index(k)
            vals[k] = v[i]
            lines.append(f"{k:75s} {v[i]:>16s} {u[i]}")
    sass = ncu_csv(rep, "source", ("--print-source", "sass"))
    hh = sass[1]
    body = sass[2:]
    ia, isrc, isamp = hh.index("Instructions Executed"), hh.index("Source"), hh.index("# Samples")
    ops, samples = Counter(), Counter()
    for r in body:
        if len(r) <= ia or not r[ia].isdigit():
            continue
        t = r[isrc].split()
        op = (t[1] if t[0].startswith("@") else t[0]).split(".")[0]
        ops[op] += int(r[ia])
        samples[op] += int(r[isamp] or 0)
    tot = sum(ops.values())
    lines += ["", f"# opcode mix (warp instructions executed, total {tot})"]
    for op, n in ops.most_common(14):
        lines.append(f"{op:10s} {n:>12d} {100.0 * n / tot:5.1f}%   stall samples {samples[op]}")
    stall_cols = [i for i, c in enumerate(hh) if c.startswith("stall_") and "Not Issued" not in c]
    st = Counter()
    for r in body:
        for i in stall_cols:
            if i < len(r) and r[i].isdigit():
                st[hh[i]] += int(r[i])
    s = sum(st.values()) or 1
    lines += ["", "# warp stall sampling (all samples)"]
    for k, n in st.most_common(8):
        lines.append(f"{k:28s} {n:>8d} {100.0 * n / s:5.1f}%")
    with open(os.path.join(PROF, f"{tag}_{name}.txt"), "w") as f:
        f.write("\n".join(lines) + "\n")
    return vals


def num(s):
    return float(s.replace(",", ""))


def main():
    tag = sys.argv[1] if len(sys.argv) > 1 else "r1"
    os.makedirs(PROF, exist_ok=True)
    for t in ("binary", "quant", "eqtl"):
        launches(tag, t)
    for name in ("k1", "fisher", "stats", "k3", "k3m", "k3m3", "k3m2x", "k4", "k4b", "general", "eqtl", "k3c", "e3c"):
        vals = full(tag, name)
        if name == "k3m" and vals:
            raw = ncu_csv(os.path.join(OUT, "prof_k3m.ncu-rep"), "raw")
            h, u, v = raw[0], raw[1], raw[2]

            def nbytes3(key):
                i = h.index(key)
                scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u[i]]
                return num(v[i]) * scale
            tr = nbytes3("dram__bytes_read.sum") + nbytes3("dram__bytes_write.sum")
            json.dump({"dram_bytes_per_launch": tr, "source": f"profiles/{tag}_k3m.txt (ncu --set full, one launch, "
                       "bench.py --trait quantitative, C3 workload)"}, open(os.path.join(PROF, "k3m_traffic.json"), "w"))
        if name == "k1" and vals:
            raw = ncu_csv(os.path.join(OUT, "prof_k1.ncu-rep"), "raw")
            h, u, v = raw[0], raw[1], raw[2]

            def nbytes(key):
                i = h.index(key)
                scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u[i]]
                return num(v[i]) * scale
            tr = nbytes("dram__bytes_read.sum") + nbytes("dram__bytes_write.sum")
            json.dump({"dram_bytes_per_launch": tr, "source": f"profiles/{tag}_k1.txt (ncu --set full, one launch, "
                       "bench.py default C3 workload)"}, open(os.path.join(PROF, "k1_traffic.json"), "w"))
    for f in ("bench_binary.json", "bench_quant.json", "bench_eqtl.json"):
        p = os.path.join(OUT, f)
        if os.path.exists(p) and os.path.getsize(p):
            with open(p) as src, open(os.path.join(PROF, f"{tag}_{f}"), "w") as dst:
                dst.write(src.read())


if __name__ == "__main__":
    main()
