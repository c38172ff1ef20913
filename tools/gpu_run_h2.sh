#!/bin/bash
# round 2h: ncu --set full of the two-tile K1 kernel (8-bit lo image, 5 ring slots) + launch list of the default bench
mkdir -p gpurun_out
timeout 200 python tools/tc_check.py f16 2 > gpurun_out/h_tc_check.log 2>&1; echo "tc_check rc=$?"
timeout 600 ncu --set full --import-source on --clock-control none -k regex:rollout_tc2_kernel -c 1 -o gpurun_out/h_tc2 python tools/tc_check.py f16 2 > gpurun_out/h_ncu_tc2.log 2>&1; echo "ncu full rc=$?"
ncu -i gpurun_out/h_tc2.ncu-rep --page raw --csv > gpurun_out/h_tc2_raw.csv 2>/dev/null
timeout 400 python bench.py --steps 2 --warmup 3 > /dev/null 2>&1; echo "bench rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/h_launches_bench.csv python bench.py --steps 2 --warmup 3 > gpurun_out/h_ncu_bench.log 2>&1; echo "ncu launches rc=$?"
rm -f gpurun_out/h_tc2.ncu-rep
