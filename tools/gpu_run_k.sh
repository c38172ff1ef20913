#!/bin/bash
# N GPUs: 2-rank NCCL pytest (fused sharded ndiv), bench at N
mkdir -p gpurun_out
N=${1:-2}
timeout 400 python -m pytest tests/test_gpu_nccl.py -m gpu -x -q > gpurun_out/k_pytest_nccl.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/k_pytest_nccl.log
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29542 tools/sharded_check.py > gpurun_out/k_sharded_n$N.log 2>&1; echo "sharded rc=$?"; grep "ndiv" gpurun_out/k_sharded_n$N.log | grep "rank 0"
timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/k_bench_n$N.json 2> gpurun_out/k_bench_n$N.err; echo "bench n$N rc=$?"
tail -c 400 gpurun_out/k_bench_n$N.err
python - <<PY
import json
d=json.loads(open('gpurun_out/k_bench_n$N.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','ms_per_step')}, d['ndiv']['ms_per_step'], d['ndiv_sharded'])
PY
