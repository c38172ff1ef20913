#!/bin/bash
mkdir -p gpurun_out
timeout 400 python bench.py --steps 20 --warmup 3 > gpurun_out/f_bench_n1.json 2> gpurun_out/f_bench_n1.err; echo "bench n1 rc=$?"
python - <<PY
import json
d=json.loads(open('gpurun_out/f_bench_n1.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','ms_per_step')}, 'e2e', d['e2e']['value'], 'roof', d['roofline']['frac'])
print('k3', {k:(round(v['ms']*1e3,1), round(v['frac_of_hbm_peak'],3)) for k,v in d['k3'].items() if isinstance(v,dict)})
print('gpu_launches', d.get('gpu_launches'), 'cpu_baseline', d['cpu_baseline']['value'], d['cpu_baseline'].get('verbatim_loop'))
PY
