#!/bin/bash
# bench with clocks: overlapped vs sequential slice path vs DMMA
mkdir -p gpurun_out
for v in "ovl:" "seq:DLA_B200_OVERLAP=0" "dmma:DLA_B200_CONTRACT=dmma"; do
  name=${v%%:*}; envs=${v#*:}
  env $envs timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$name.json 2> gpurun_out/bench_$name.err; echo "bench $name rc=$?"
  python - <<PY
import json
d=json.load(open("gpurun_out/bench_$name.json"))
print("$name", "value", round(d["value"],1), "e2e", round(d["e2e"]["value"],1), "ms/step", round(d["ms_per_step"],2), d["clocks"], {k: round(v,2) for k,v in d["kernel_ms_per_step"].items()})
PY
done
