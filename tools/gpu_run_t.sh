#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/encoder_tc_check.py > gpurun_out/t_enc_check.txt 2>&1; echo "check rc=$?"; grep -v "Warn\|warn" gpurun_out/t_enc_check.txt | tail -14 | cut -c1-170
timeout 300 python -m pytest tests/test_gpu_encoder.py -m gpu -x -q > gpurun_out/t_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/t_pytest.log
timeout 200 python tools/encoder_time.py > gpurun_out/t_enc_time.txt 2>&1; echo "time rc=$?"; tail -5 gpurun_out/t_enc_time.txt
