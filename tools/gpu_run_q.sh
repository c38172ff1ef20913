#!/bin/bash
# 2 GPUs: whole GPU suite (incl. the 2-rank NCCL test with the skewed ring), smoke, bench N=1 and N=2
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/g_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/g_pytest_gpu.log
grep -h "skewed ring" gpurun_out/g_pytest_gpu.log | head -3
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/g_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/g_smoke.log
timeout 400 python bench.py --steps 20 --warmup 3 > gpurun_out/g_bench_n1.json 2> gpurun_out/g_bench_n1.err; echo "bench n1 rc=$?"
timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/g_bench_n2.json 2> gpurun_out/g_bench_n2.err; echo "bench n2 rc=$?"
python - <<PY
import json
for n in (1, 2):
    d=json.loads(open('gpurun_out/g_bench_n%d.json' % n).read().strip().splitlines()[-1])
    s=dict(d['ndiv_sharded']); s.pop('how',None); s.pop('timing',None)
    print(n, {k:d[k] for k in ('value','ms_per_step')}, 'e2e', d['e2e']['value'], 'roof', d['roofline']['frac'], 'k3', d.get('k3',{}).get('argmin'), 'ndiv', d['ndiv']['ms_per_step'], s)
PY
