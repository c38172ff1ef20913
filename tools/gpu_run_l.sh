#!/bin/bash
# 1 GPU: select kernels after the float-domain argmin change
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_select.py tests/test_gpu_mpc.py -m gpu -x -q > gpurun_out/l_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/l_pytest.log
timeout 200 python tools/select_time.py > gpurun_out/l_select_time.txt 2>&1; echo "select_time rc=$?"; cat gpurun_out/l_select_time.txt
timeout 300 ncu --set full --clock-control none --import-source on -k regex:argmin_rows_kernel -s 5 -c 1 -o gpurun_out/l_ncu_argmin env NDP_SELECT_ONLY=4096x4096 python tools/select_time.py > gpurun_out/l_ncu_argmin.log 2>&1; echo "ncu rc=$?"
