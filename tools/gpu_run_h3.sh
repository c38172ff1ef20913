#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_mpc.py -x -q -k "closed_loop_sweep_matches" 2>&1 | tail -3
timeout 400 python bench.py --steps 20 --warmup 3 > gpurun_out/h3_bench_n1.json 2> gpurun_out/h3_bench_n1.err; echo "bench n1 rc=$?"; tail -3 gpurun_out/h3_bench_n1.err
python - <<PY
import json
d=json.loads(open('gpurun_out/h3_bench_n1.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','ms_per_step')}, 'e2e', d['e2e']['value'], 'roof', d['roofline']['frac'], d['roofline']['traffic'])
print(d['modes']['E256'])
PY
