#!/bin/bash
# 1 GPU: whole GPU suite, two-tile kernel check, smoke, bench N=1 (round 2h build)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/h_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/h_pytest_gpu.log
timeout 240 python tools/tc2_check.py > gpurun_out/h_tc2_check.txt 2>&1; echo "tc2_check rc=$?"; tail -3 gpurun_out/h_tc2_check.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/h_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/h_smoke.log
timeout 400 python bench.py --steps 20 --warmup 3 > gpurun_out/h_bench_n1.json 2> gpurun_out/h_bench_n1.err; echo "bench n1 rc=$?"
python - <<PY
import json
d=json.loads(open('gpurun_out/h_bench_n1.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','ms_per_step')}, 'e2e', d['e2e']['value'], 'roof', d['roofline']['frac'], d['roofline']['kernel_ms'], 'sus', d['sustained']['candidate_steps_per_s'], d['sustained']['frac_of_sustained_peak'])
print('verify', {k:d['verify'][k] for k in ('plans','verified1','fallback2')}, 'parity', d['parity']['cost_rel_err_all_max'], d['parity']['argmin_equal'], d['parity']['topk_set_equal'])
print('cfg4', d.get('cfg4',{}).get('value'), 'cfg5', d.get('cfg5',{}).get('value'), 'ndiv', d['ndiv']['ms_per_step'])
PY
