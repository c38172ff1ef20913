#!/bin/bash
mkdir -p gpurun_out
timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/h_bench_n2.json 2> gpurun_out/h_bench_n2.err; echo "bench n2 rc=$?"
tail -c 800 gpurun_out/h_bench_n2.err; head -c 600 gpurun_out/h_bench_n2.json
