#!/bin/bash
mkdir -p gpurun_out
timeout 300 ncu --set full --clock-control none --import-source on -k regex:conv_tc_kernel -s 8 -c 4 -o gpurun_out/y_ncu_conv_tc python tools/encoder_once.py 64 > gpurun_out/y_ncu.log 2>&1; echo "ncu rc=$?"; tail -3 gpurun_out/y_ncu.log
