#!/bin/bash
mkdir -p gpurun_out
for lib in libndivplan_prof.so libndivplan_exp5.so; do
echo "== $lib"
NDP_TC2_NOSPLIT=1 NDP_LIB_PATH=ndivplanning_b200/$lib timeout 200 python tools/tc2_prof.py 2>&1 | head -4
done
