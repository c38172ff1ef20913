#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/tc2_mini.py > gpurun_out/r2_mini.txt 2>&1; echo "rc=$?"; tail -12 gpurun_out/r2_mini.txt | cut -c1-220
