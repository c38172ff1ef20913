#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_encoder.py -m gpu -x -q > gpurun_out/x_pytest.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/x_pytest.log | cut -c1-220
