#!/bin/bash
mkdir -p gpurun_out
timeout 240 python tools/tc2_check.py > gpurun_out/r_tc2_check.txt 2>&1; echo "tc2_check rc=$?"; grep -v "^cfg" gpurun_out/r_tc2_check.txt | tail -8
timeout 600 python -m pytest tests/test_gpu_mpc.py tests/test_gpu_baseline_cfgs.py -m gpu -x -q > gpurun_out/r_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r_pytest.log
