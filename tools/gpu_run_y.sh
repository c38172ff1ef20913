#!/bin/bash
mkdir -p gpurun_out
timeout 300 ncu --set full --clock-control none --import-source on -k regex:conv1_kernel -s 1 -c 1 -o gpurun_out/y_ncu_conv1 python tools/encoder_once.py 64 > gpurun_out/y_ncu.log 2>&1; echo "ncu rc=$?"; tail -2 gpurun_out/y_ncu.log
