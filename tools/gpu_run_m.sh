#!/bin/bash
mkdir -p gpurun_out
N=${1:-2}
for n in 1 $N; do
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29547 tools/sharded_ndiv_trace.py > gpurun_out/m_trace_n$n.log 2>&1; echo "trace n$n rc=$?"; grep -A14 "^world" gpurun_out/m_trace_n$n.log
done
