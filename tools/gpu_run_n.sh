#!/bin/bash
# N GPUs: peer-window correctness at N ranks, per-kernel trace of the sharded ndiv step, bench
mkdir -p gpurun_out
N=${1:-8}
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29542 tools/sharded_check.py > gpurun_out/n_sharded_n$N.log 2>&1; echo "sharded rc=$?"; grep "rank 0" gpurun_out/n_sharded_n$N.log | cut -c1-250
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29547 tools/sharded_ndiv_trace.py > gpurun_out/n_trace_n$N.log 2>&1; echo "trace n$N rc=$?"; grep -A10 "^world" gpurun_out/n_trace_n$N.log | cut -c1-160
timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/n_bench_n$N.json 2> gpurun_out/n_bench_n$N.err; echo "bench n$N rc=$?"
tail -c 300 gpurun_out/n_bench_n$N.err
python - <<PY
import json
d=json.loads(open('gpurun_out/n_bench_n$N.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','ms_per_step')}, 'e2e', d['e2e']['value'], 'cfg4', d['cfg4'].get('value'), 'cfg5', d['cfg5'].get('value'))
s=d['ndiv_sharded']; s.pop('how',None); print(d['ndiv']['ms_per_step'], s)
PY
