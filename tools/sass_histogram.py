#!/usr/bin/env python
"""Opcode histogram of the sm_100a SASS of libdla_b200.so, per kernel (cuobjdump -sass): the evidence that the slice
path issues tcgen05 MMAs (UTCIMMA = kind::i8, UTCHMMA/UTCQMMA others), bulk async copies (UBLKCP), TMEM loads (LDTM),
mbarrier waits (SYNCS) and FP64 FMAs (DFMA).      python tools/sass_histogram.py > profiles/r02_sass_opcodes.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "gp_dla_detection_b200", "libdla_b200.so")
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
WATCH = ["UTCIMMA", "UTCHMMA", "UTCQMMA", "UTCBAR", "UTCATOMSWS", "UBLKCP", "UTMALDG", "LDTM", "STTM", "SYNCS", "ELECT", "DMMA",
         "DFMA", "DMUL", "DADD", "MUFU", "IMAD", "PRMT", "I2F", "F2F", "LDS", "STS", "LDG", "STG", "LDL", "STL", "BAR", "R2UR"]
kern, hist = None, collections.OrderedDict()
for line in sass.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        kern = m.group(1)
        hist[kern] = collections.Counter()
        continue
    m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
    if m and kern:
        hist[kern][m.group(1)] += 1
demangle = subprocess.run(["c++filt"], input="\n".join(hist), capture_output=True, text=True).stdout.splitlines()
print("cuobjdump -sass gp_dla_detection_b200/libdla_b200.so (sm_100a), opcode counts per kernel")
print("columns: total instructions, then the watched opcodes that occur (mnemonic before the first '.')\n")
want = sys.argv[1:] or None
for (k, h), name in zip(hist.items(), demangle):
    if want and not any(w in name for w in want):
        continue
    short = re.sub(r"\(.*", "", name)
    tot = sum(h.values())
    parts = [f"{op}={h[op]}" for op in WATCH if h.get(op)]
    print(f"{short}\n    total={tot}  " + "  ".join(parts))
