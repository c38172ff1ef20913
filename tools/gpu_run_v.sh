#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/v_pytest_alone.log 2>&1; echo "alone rc=$?"; tail -3 gpurun_out/v_pytest_alone.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/v_pytest_all.log 2>&1; echo "all rc=$?"; tail -3 gpurun_out/v_pytest_all.log
