#!/bin/bash
mkdir -p gpurun_out
N=${1:-8}
timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/g_bench_n$N.json 2> gpurun_out/g_bench_n$N.err; echo "bench n$N rc=$?"
tail -c 300 gpurun_out/g_bench_n$N.err
python - <<PY
import json
d=json.loads(open('gpurun_out/g_bench_n$N.json').read().strip().splitlines()[-1])
s=dict(d['ndiv_sharded']); s.pop('how',None); s.pop('timing',None)
print({k:d[k] for k in ('value','ms_per_step')}, 'e2e', d['e2e']['value'], 'cfg4', d['cfg4']['candidate_steps_per_s'], 'cfg5', d['cfg5']['candidate_steps_per_s'], 'sus', d['sustained']['candidate_steps_per_s'], 'roof', d['roofline']['frac'])
print(d['ndiv']['ms_per_step'], d['ndiv'].get('value_all_ranks'), s, d['verify']['verified1'], d['clocks'])
PY
