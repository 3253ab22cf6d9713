#!/usr/bin/env python
"""Executed-instruction mix and stall samples per opcode from an ncu source page exported as CSV:
    ncu -i X.ncu-rep --page source --csv --print-source sass > x.csv ;  python tools/ncu_sass_mix.py x.csv [pairs]
pairs = (hypothesis, pixel) pairs of the launch: prints warp instructions x 32 / pairs per opcode."""
import collections
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
pairs = float(sys.argv[2]) if len(sys.argv) > 2 else None
hdr = rows[1]
ia, isrc, iex, ismp = hdr.index("Address"), hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
stall_cols = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
ex, smp, stalls = collections.Counter(), collections.Counter(), collections.Counter()
tot_ex = tot_smp = 0
hot = []
for r in rows[2:]:
    if len(r) <= iex or not r[iex]:
        continue
    m = re.match(r"\s*(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", r[isrc])
    if not m:
        continue
    op = m.group(1)
    n, s = int(r[iex]), int(r[ismp] or 0)
    ex[op] += n
    smp[op] += s
    tot_ex += n
    tot_smp += s
    for i, h in stall_cols:
        if r[i]:
            stalls[h] += int(r[i])
    hot.append((s, n, r[ia], r[isrc][:90]))
print(f"warp instructions executed {tot_ex}, stall samples {tot_smp}")
print(f"{'opcode':10s} {'executed':>12s} {'share':>7s} {'per pair':>9s} {'samples':>9s} {'share':>7s}")
for op, n in ex.most_common(28):
    pp = f"{32.0 * n / pairs:9.2f}" if pairs else ""
    print(f"{op:10s} {n:12d} {100.0 * n / tot_ex:6.2f}% {pp} {smp[op]:9d} {100.0 * smp[op] / max(1, tot_smp):6.2f}%")
print("\nstall reasons (samples):")
for h, n in stalls.most_common(12):
    print(f"  {h:24s} {n:9d} {100.0 * n / max(1, tot_smp):6.2f}%")
print("\nhottest instructions by samples:")
for s, n, a, src in sorted(hot, reverse=True)[:40]:
    print(f"  {s:7d} samples  {n:10d} exec  {a}  {src}")
