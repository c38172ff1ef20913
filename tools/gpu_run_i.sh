#!/bin/bash
mkdir -p gpurun_out
N=${1:-8}
timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/i_bench_n$N.json 2> gpurun_out/i_bench_n$N.err; echo "bench n$N rc=$?"
tail -c 600 gpurun_out/i_bench_n$N.err; head -c 300 gpurun_out/i_bench_n$N.json
