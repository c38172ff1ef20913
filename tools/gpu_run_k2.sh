#!/bin/bash
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_nccl.py tests/test_gpu_select.py tests/test_gpu_ndiv.py -m gpu -x -q > gpurun_out/k_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/k_pytest.log
