#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/tc2_check.py > gpurun_out/r_tc2_check.txt 2>&1; echo "tc2_check rc=$?"; cat gpurun_out/r_tc2_check.txt | tail -20
timeout 600 python -m pytest tests/test_gpu_mpc.py tests/test_gpu_baseline_cfgs.py -m gpu -x -q > gpurun_out/r_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r_pytest.log
timeout 400 python bench.py --steps 20 --warmup 3 > gpurun_out/r_bench_n1.json 2> gpurun_out/r_bench_n1.err; echo "bench n1 rc=$?"
NDP_TC2_NOSPLIT=1 timeout 400 python bench.py --steps 20 --warmup 3 > gpurun_out/r_bench_n1_nosplit.json 2> gpurun_out/r_bench_n1_nosplit.err; echo "bench nosplit rc=$?"
python - <<PY
import json
for f in ('r_bench_n1','r_bench_n1_nosplit'):
    d=json.loads(open('gpurun_out/%s.json' % f).read().strip().splitlines()[-1])
    print(f, {k:d[k] for k in ('value','ms_per_step')}, 'e2e', d['e2e']['value'], 'roof', d['roofline']['frac'], d['roofline'].get('kernel_ms'), 'single', d.get('single_plan_ms'), 'cfg4', d['cfg4']['candidate_steps_per_s'], 'cfg5', d['cfg5']['candidate_steps_per_s'], 'sus', d['sustained']['candidate_steps_per_s'])
PY
