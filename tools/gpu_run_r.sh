#!/bin/bash
mkdir -p gpurun_out
timeout 240 python tools/tc2_check.py > gpurun_out/r_tc2_check.txt 2>&1; echo "tc2_check rc=$?"; tail -22 gpurun_out/r_tc2_check.txt | cut -c1-200
NDP_TC2_NOSPLIT=1 NDP_LIB_PATH=ndivplanning_b200/libndivplan_prof.so timeout 120 python tools/tc2_prof.py 2>&1 | head -4 > gpurun_out/r_tc2_prof_cluster.txt; cat gpurun_out/r_tc2_prof_cluster.txt
