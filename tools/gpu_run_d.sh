#!/bin/bash
# round-2 GPU call D: two-tile K1 check + timing, MPC tests, select timing, ncu of the ndiv kernels
mkdir -p gpurun_out
timeout 300 python tools/tc2_check.py > gpurun_out/d_tc2_check.txt 2>&1; echo "tc2_check rc=$?" >> gpurun_out/d_tc2_check.txt
cat gpurun_out/d_tc2_check.txt | tail -12
timeout 1500 python -m pytest tests/test_gpu_mpc.py tests/test_gpu_baseline_cfgs.py tests/test_gpu_select.py -m gpu -q -x --timeout=600 > gpurun_out/d_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/d_pytest.log
tail -6 gpurun_out/d_pytest.log
timeout 300 python tools/select_time.py > gpurun_out/d_select_time.txt 2>&1; cat gpurun_out/d_select_time.txt
timeout 600 ncu --set full --import-source on --clock-control none -k regex:ndiv_gram_kernel -c 4 -o gpurun_out/d_ndiv_gram python tools/ndiv_time.py cfg3 > gpurun_out/d_ncu_ndiv.log 2>&1
ncu -i gpurun_out/d_ndiv_gram.ncu-rep --page raw --csv > gpurun_out/d_ndiv_gram_raw.csv 2>/dev/null
ls -la gpurun_out/d_*
