#!/usr/bin/env python
"""Summarise gpurun_out/r02_slices.ncu-rep (ncu --set full of the slice-path kernels) and gpurun_out/g_launches.csv
(ncu --metrics gpu__time_duration.sum of bench.py --steps 2 --warmup 3) into profiles/r02_ncu_slices_summary.txt and
profiles/r02_ncu_launches.txt."""
import collections
import csv
import io
import json
import subprocess

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.sum", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tensor.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_subpipe_imma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_uniform.sum", "sm__inst_executed_pipe_alu.sum", "sm__inst_executed_pipe_fma.sum",
        "sm__inst_executed_pipe_lsu.sum", "sm__inst_executed_pipe_xu.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct", "sm__cycles_elapsed.max", "sm__cycles_active.avg",
        "smsp__cycles_active.avg", "sm__throughput.avg.pct_of_peak_sustained_elapsed"]

raw = subprocess.run(["ncu", "-i", "gpurun_out/r02_slices.ncu-rep", "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
seen = set()
facts = {}
with open("profiles/r02_ncu_slices_summary.txt", "w") as f:
    f.write("ncu --set full --clock-control none --import-source on, python bench.py --steps 1 --warmup 1 --quasars 16 --no-cpu-baseline\n"
            "(per launch: one chunk of 16 quasars x 10000 hypotheses (S_pad 10112), 1217 px (n_pad32 1248, 39 K steps), k=20;\n"
            " profiler-serialised, cold cache; first captured launch of each kernel)\n\n")
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        name = d["Kernel Name"].split("(")[0]
        if name in seen:
            continue
        seen.add(name)
        f.write(f"== {name}\n")
        for k in KEYS:
            if k in d and d[k] not in ("", "n/a"):
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
        def num(k):
            try:
                return float(d.get(k, "nan").replace(",", ""))
            except ValueError:
                return float("nan")
        facts[name] = {"dram_read": num("dram__bytes_read.sum"), "dram_read_unit": units[hdr.index("dram__bytes_read.sum")],
                       "dram_write": num("dram__bytes_write.sum"), "dram_write_unit": units[hdr.index("dram__bytes_write.sum")],
                       "fp64_inst": num("sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active") / 100.0 * num("sm__cycles_active.avg") * 2.0 * 148,
                       "tensor_pct": num("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"), "time": num("gpu__time_duration.sum"),
                       "time_unit": units[hdr.index("gpu__time_duration.sum")], "inst": num("smsp__inst_executed.sum")}
    f.write("derived (16 quasars, S_pad = 10112, n_pad32 = 1248 -> 201 916 416 (hypothesis, pixel) pairs):\n")
    pairs = 16 * 10112 * 1248.0
    for name, v in facts.items():
        if "voigt_slices" in name:
            f.write("  (FP64 warp instructions = pipe_fp64 pct_of_peak_sustained_active x sm__cycles_active.avg x 2 per clock per SM x 148 SMs:\n"
                    "   the FP64 pipe of an SM retires 64 lanes = 2 warp instructions per clock, which is what cuBLAS DGEMM's 35-37 TFLOP/s implies)\n")
            f.write(f"  K1s: warp-level FP64 instructions {v['fp64_inst']:.0f} x 32 lanes / pairs = {v['fp64_inst'] * 32 / pairs:.2f} per (hypothesis, pixel); "
                    f"all instructions {v['inst'] * 32 / pairs:.1f} per pair\n")
    for name, v in facts.items():
        if "contract_slices" in name:
            ops = 2.0 * 27 * 256 * pairs
            f.write(f"  K2s: int8 operations 2 x 27 slice pairs x 256 columns x pairs = {ops:.4g}; / {v['time']} ms = {ops / v['time'] / 1e9:.0f} TOP/s; "
                    f"tensor pipe active {v['tensor_pct']:.1f} % of cycles; DRAM {v['dram_read'] + v['dram_write']:.3f} GB per launch against "
                    f"algorithmic planes {12 * pairs / 1e9:.3f} GB + tables + G {16 * 10112 * 256 * 8 / 1e9:.3f} GB\n")
    f.write(json.dumps(facts) + "\n")
print(open("profiles/r02_ncu_slices_summary.txt").read())

# ---- launch list: per-kernel totals and shares of the step -------------------------------------------------------
tot, n = collections.OrderedDict(), collections.Counter()
lines = [l for l in open("gpurun_out/g_launches.csv") if l.startswith('"')]
rd = csv.reader(lines)
h = next(rd)
ik, iv, iu = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
for r in rd:
    v = float(r[iv].replace(",", ""))
    v *= {"ns": 1e-6, "us": 1e-3, "usecond": 1e-3, "ms": 1.0, "msecond": 1.0, "nsecond": 1e-6, "s": 1e3, "second": 1e3}.get(r[iu], 1.0)
    name = r[ik].split("(")[0]
    tot[name] = tot.get(name, 0.0) + v
    n[name] += 1
ours = lambda k: "dla::" in k or "sl::" in k or k.endswith("_kernel")
bench = json.load(open("gpurun_out/g_bench_short.json"))
km = bench["kernel_ms_per_step"]
with open("profiles/r02_ncu_launches.txt", "w") as f:
    f.write("ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv, python bench.py --steps 2 --warmup 3 --no-cpu-baseline\n"
            "(the first 400 launches: the DGEMM peak measurement, then the warm-up and timed steps of the device-resident leg).\n"
            "Times are cold-cache and serialised: compare SHARES with bench.py's kernel_ms_per_step, not absolutes.\n\n")
    sd = sum(v for k, v in tot.items() if ours(k))
    f.write("kernels of the step, share of the step under ncu:\n")
    for k, v in sorted(tot.items(), key=lambda kv: -kv[1]):
        if ours(k):
            f.write(f"  {k:58s} launches {n[k]:4d}  total {v:10.3f} ms  share {100 * v / sd:6.2f} %\n")
    f.write("\nsame command without ncu (gpurun_out/g_bench_short.json, CUDA events on the library's stream), ms per step and share:\n")
    st = km["ms_total"]
    for k in ("ms_phi", "ms_voigt", "ms_contract", "ms_chol", "ms_reduce"):
        f.write(f"  {k:12s} {km[k]:8.3f} ms  {100 * km[k] / st:6.2f} %\n")
    f.write(f"  ms_total     {st:8.3f} ms   value {bench['value']:.0f} quasars/s\n")
    f.write("\nother launches of the same command (cuBLAS DGEMM = the live FP64 peak measurement; torch RNG for it):\n")
    for k, v in sorted(tot.items(), key=lambda kv: -kv[1]):
        if not ours(k):
            f.write(f"  {k[:90]:90s} launches {n[k]:4d}  total {v:10.3f} ms\n")
print(open("profiles/r02_ncu_launches.txt").read())
