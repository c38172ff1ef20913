#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_mpc.py tests/test_gpu_baseline_cfgs.py tests/test_gpu_encoder.py -m gpu -x -q > gpurun_out/w_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/w_pytest.log | cut -c1-220
