#!/bin/bash
# round-2 evidence on one GPU: tests, bench lines, ncu launch list + full capture, parity records
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/f_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/f_pytest.log
timeout 900 python bench.py > gpurun_out/h_bench_default.json 2> gpurun_out/h_bench_default.err; echo "default rc=$?"
for cfg in "config3 64" "config4 8" "config5 16"; do
  set -- $cfg
  timeout 600 python bench.py --workload $1 --quasars $2 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/h_cfg_$1.json 2> gpurun_out/h_cfg_$1.err; echo "$1 rc=$?"
done
DLA_B200_CONTRACT=dmma timeout 600 python bench.py --workload config5 --quasars 16 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/h_cfg_config5_dmma.json 2>/dev/null; echo "config5 dmma rc=$?"
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/h_bench_reference.json 2> gpurun_out/h_bench_reference.err; echo "reference rc=$?"
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$B > gpurun_out/g_bench_short.json 2> gpurun_out/g_bench_short.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/g_launches.csv $B > gpurun_out/g_ncu_launch.log 2>&1
echo "launch list rc=$?"
F="python bench.py --steps 1 --warmup 1 --quasars 16 --no-cpu-baseline"
$F > gpurun_out/g_bench16.json 2> gpurun_out/g_bench16.err && \
timeout 1200 ncu --set full --clock-control none --import-source on \
    -k regex:'voigt_slices|contract_slices|chol_rows|table_slices|table_colscale|pixel_scales|reduce|prep_model' -c 9 \
    -f -o gpurun_out/r02_slices $F > gpurun_out/g_ncu_full.log 2>&1
echo "ncu full rc=$?"
timeout 600 python tools/parity_truth.py > gpurun_out/k_truth.log 2>&1; echo "truth rc=$?"
timeout 1400 python tools/parity_full.py > gpurun_out/parity_full.log 2>&1; echo "parity_full rc=$?"
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/h_*.json')):
    try:
        d=json.loads([l for l in open(f) if l.startswith('{')][-1])
        k=d.get('kernel_ms_per_step',{})
        print(f, 'value %.1f e2e %.1f'%(d['value'], d['e2e']['value']), 'roof', d.get('roofline',{}).get('frac'), {a: round(b,3) for a,b in k.items()})
    except Exception as e:
        print(f, 'failed', e)
PY
