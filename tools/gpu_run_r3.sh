#!/bin/bash
mkdir -p gpurun_out
timeout 240 python tools/tc2_check.py > gpurun_out/r_tc2_check.txt 2>&1; echo "tc2_check rc=$?"; grep -v "^cfg" gpurun_out/r_tc2_check.txt | tail -8
timeout 240 python tools/tc2_check.py 2>&1 | grep "rollout:"
