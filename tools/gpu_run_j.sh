#!/bin/bash
# 1 GPU: fused sharded-ndiv entry (world-1), ncu evidence for pass 2 / select kernels
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_ndiv.py tests/test_gpu_select.py -m gpu -x -q > gpurun_out/j_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/j_pytest.log
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 1 --master-addr 127.0.0.1 --master-port 29541 tools/sharded_check.py > gpurun_out/j_sharded_n1.log 2>&1; echo "sharded n1 rc=$?"; grep ndiv gpurun_out/j_sharded_n1.log
timeout 300 python bench.py --steps 10 --warmup 3 > gpurun_out/j_bench_n1.json 2> gpurun_out/j_bench_n1.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/j_bench_n1.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','ms_per_step')}, d['ndiv']['ms_per_step'], d['ndiv_sharded'])
PY
timeout 300 ncu --set full --clock-control none --import-source on -k regex:ndiv_gram_kernel -s 6 -c 2 -o gpurun_out/j_ncu_ndiv python tools/ndiv_time.py > gpurun_out/j_ncu_ndiv.log 2>&1; echo "ncu ndiv rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:"select2_kernel|argmin_rows_kernel" -s 5 -c 2 -o gpurun_out/j_ncu_argmin env NDP_SELECT_ONLY=4096x4096 python tools/select_time.py > gpurun_out/j_ncu_select.log 2>&1; echo "ncu select rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:select2_kernel -s 5 -c 2 -o gpurun_out/j_ncu_top16 env NDP_SELECT_ONLY=4096x4096 python tools/select_time.py > gpurun_out/j_ncu_top16.log 2>&1; echo "ncu top16 rc=$?"
ls -la gpurun_out | tail -8
