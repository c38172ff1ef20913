#!/bin/bash
mkdir -p gpurun_out
timeout 300 python bench.py --steps 20 --warmup 3 > gpurun_out/h4_bench_n1.json 2> gpurun_out/h4_bench_n1.err; echo "bench n1 rc=$?"
python - <<PY
import json
d=json.loads(open('gpurun_out/h4_bench_n1.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','ms_per_step')}, 'e2e', d['e2e']['value'], 'roof', d['roofline']['frac'], d['roofline']['kernel_ms'], 'sus', d['sustained']['candidate_steps_per_s'], d['sustained']['frac_of_sustained_peak'])
print('verify', {k:d['verify'][k] for k in ('plans','verified1','fallback2')}, 'parity', d['parity']['cost_rel_err_all_max'], d['parity']['argmin_equal'], d['parity']['topk_set_equal'])
PY
timeout 60 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
