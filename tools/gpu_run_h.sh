#!/bin/bash
# round-2 GPU call H (2 GPUs): NCCL pytest (sharded plan + ShardedNdiv), select tests, bench at N=2
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/h_gpus.txt
timeout 900 python -m pytest tests/test_gpu_nccl.py tests/test_gpu_select.py -m gpu -q --timeout=600 > gpurun_out/h_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/h_pytest.log
tail -15 gpurun_out/h_pytest.log
timeout 300 python tools/select_time.py > gpurun_out/h_select_time.txt 2>&1; tail -3 gpurun_out/h_select_time.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/h_bench_n2.json 2> gpurun_out/h_bench_n2.err; echo "bench n2 rc=$?"
tail -c 1500 gpurun_out/h_bench_n2.err
