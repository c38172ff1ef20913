#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_mpc.py -m gpu -x -q > gpurun_out/s_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/s_pytest.log
timeout 400 python bench.py --steps 20 --warmup 3 > gpurun_out/s_bench_n1.json 2> gpurun_out/s_bench_n1.err; echo "bench n1 rc=$?"
python - <<PY
import json
for f in ('s_bench_n1',):
    d=json.loads(open('gpurun_out/%s.json' % f).read().strip().splitlines()[-1])
    print(f, {k:d[k] for k in ('value','ms_per_step')}, 'e2e', d['e2e']['value'], 'roof', d['roofline']['frac'], d['roofline'].get('kernel_ms'), 'cfg4', d['cfg4']['candidate_steps_per_s'], 'cfg5', d['cfg5']['candidate_steps_per_s'], 'sus', d['sustained']['candidate_steps_per_s'])
    print({k:v for k,v in d.items() if 'single' in k or 'latency' in k})
PY
