#!/bin/bash
mkdir -p gpurun_out
NDP_LIB_PATH=ndivplanning_b200/libndivplan_prof.so timeout 200 python tools/gram_prof.py > gpurun_out/p_gram_prof.txt 2>&1; echo "rc=$?"; cat gpurun_out/p_gram_prof.txt
