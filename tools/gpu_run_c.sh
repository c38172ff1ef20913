#!/bin/bash
# round-2 GPU call C: probe (MMA issue floor), ndiv + select tests, ndiv timing + launch list, select timing
mkdir -p gpurun_out
timeout 120 ./tools/tmem_probe > gpurun_out/c_tmem_probe.txt 2>&1; echo "probe rc=$?" >> gpurun_out/c_tmem_probe.txt
timeout 1200 python -m pytest tests/test_gpu_ndiv.py tests/test_gpu_select.py tests/test_gpu_baseline_cfgs.py tests/test_gpu_mpc.py -m gpu -q --timeout=900 > gpurun_out/c_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c_pytest.log
tail -8 gpurun_out/c_pytest.log
timeout 300 python tools/ndiv_time.py > gpurun_out/c_ndiv_time.txt 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/c_launches_ndiv_cfg3.csv python tools/ndiv_time.py cfg3 > /dev/null 2>&1
timeout 300 python tools/select_time.py > gpurun_out/c_select_time.txt 2>&1
cat gpurun_out/c_ndiv_time.txt gpurun_out/c_select_time.txt
