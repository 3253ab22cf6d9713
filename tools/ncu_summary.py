#!/usr/bin/env python
"""Summarise gpurun_out/prof.ncu-rep (ncu --set full) and gpurun_out/launches.csv
(ncu --metrics gpu__time_duration.sum) into small text files under profiles/."""
import csv
import io
import subprocess
import sys
import collections

tag = sys.argv[1] if len(sys.argv) > 1 else "r01"
KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tensor_subpipe_dmma.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_subpipe_dmma_cycles_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "lts__t_bytes.sum", "sm__cycles_elapsed.max"]

raw = subprocess.run(["ncu", "-i", "gpurun_out/prof.ncu-rep", "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
with open(f"profiles/{tag}_ncu_full_summary.txt", "w") as f:
    f.write("ncu --set full --clock-control none, bench.py --steps 1 --warmup 1 --quasars 16 --no-cpu-baseline\n"
            "(per launch: one chunk of 16 quasars x 10000 hypotheses, 1217 px, k=20; profiler-serialised, cold cache)\n\n")
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        f.write(f"== {d['Kernel Name'][:60]}\n")
        for k in KEYS:
            if k in d:
                f.write(f"  {k:88s} {d[k]:>18s} {units[hdr.index(k)]}\n")
        for h in hdr:
            if "average_warps_issue_stalled" in h and h.endswith("per_issue_active.ratio"):
                try:
                    v = float(d[h])
                except ValueError:
                    continue
                if v > 0.1:
                    f.write(f"  stall {h.split('stalled_')[1].split('_per_issue')[0]:40s} {v:8.3f} warps/issue\n")
        f.write("\n")

# launch list: per-kernel totals and shares
tot = collections.OrderedDict()
n = collections.Counter()
lines = [l for l in open("gpurun_out/launches.csv") if l.startswith('"')]
rd = csv.reader(lines)
h = next(rd)
ik, iv, iu = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
for r in rd:
    v = float(r[iv].replace(",", ""))
    v *= {"ns": 1e-6, "us": 1e-3, "usecond": 1e-3, "ms": 1.0, "msecond": 1.0, "nsecond": 1e-6, "s": 1e3, "second": 1e3}.get(r[iu], 1.0)
    name = r[ik].split("(")[0]
    tot[name] = tot.get(name, 0.0) + v
    n[name] += 1
s = sum(tot.values())
with open(f"profiles/{tag}_ncu_launches.txt", "w") as f:
    f.write("ncu --metrics gpu__time_duration.sum --clock-control none -c 200, same command; times are cold-cache and\n"
            "serialised: compare SHARES with bench.py's kernel_ms_per_step, not absolutes\n\n")
    sd = sum(v for k, v in tot.items() if k.startswith("dla::"))
    f.write("kernels of the timed step (dla::*), share of the step:\n")
    for k, v in sorted(tot.items(), key=lambda kv: -kv[1]):
        if k.startswith("dla::"):
            f.write(f"  {k:46s} launches {n[k]:4d}  total {v:10.3f} ms  share {100 * v / sd:6.2f} %\n")
    f.write("\nother launches of the same command (cuBLAS DGEMM = the live FP64 peak measurement; torch RNG for it):\n")
    for k, v in sorted(tot.items(), key=lambda kv: -kv[1]):
        if not k.startswith("dla::"):
            f.write(f"  {k[:90]:90s} launches {n[k]:4d}  total {v:10.3f} ms\n")
print(open(f"profiles/{tag}_ncu_launches.txt").read())
