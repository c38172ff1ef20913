#!/bin/bash
# round-2 GPU call F: first real run of the two-tile K1 kernel (bit-equality vs one-tile), counters, MPC tests
mkdir -p gpurun_out
timeout 300 python tools/tc2_check.py > gpurun_out/f_tc2_check.txt 2>&1; echo "tc2_check rc=$?" >> gpurun_out/f_tc2_check.txt
tail -12 gpurun_out/f_tc2_check.txt
NDP_LIB_PATH=$PWD/ndivplanning_b200/libndivplan_prof.so timeout 300 python tools/tc2_prof.py > gpurun_out/f_tc2_prof.txt 2>&1; echo "rc=$?" >> gpurun_out/f_tc2_prof.txt
cat gpurun_out/f_tc2_prof.txt
timeout 1500 python -m pytest tests/test_gpu_mpc.py tests/test_gpu_baseline_cfgs.py -m gpu -q -x --timeout=600 > gpurun_out/f_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/f_pytest.log
tail -6 gpurun_out/f_pytest.log
