#!/bin/bash
mkdir -p gpurun_out
N=${1:-2}
timeout 300 python -m pytest tests/test_gpu_ndiv.py tests/test_gpu_baseline_cfgs.py -m gpu -x -q -k ndiv > gpurun_out/o_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/o_pytest.log
timeout 200 python tools/ndiv_time.py > gpurun_out/o_ndiv_time.txt 2>&1; echo "ndiv_time rc=$?"; cat gpurun_out/o_ndiv_time.txt
for n in 1 $N; do
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29547 tools/sharded_ndiv_trace.py > gpurun_out/o_trace_n$n.log 2>&1; echo "trace n$n rc=$?"; grep -A8 "^world" gpurun_out/o_trace_n$n.log | cut -c1-120
done
