#!/bin/bash
mkdir -p gpurun_out
NDP_LIB_PATH=$PWD/ndivplanning_b200/libndivplan_prof.so timeout 300 python tools/tc2_prof.py > gpurun_out/e_tc2_prof.txt 2>&1; echo "rc=$?" >> gpurun_out/e_tc2_prof.txt
cat gpurun_out/e_tc2_prof.txt
