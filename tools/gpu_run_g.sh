#!/bin/bash
# round-2 GPU call G: whole GPU suite, smoke, default bench, ncu launch list of the bench, ncu --set full of the two-tile kernel
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q --timeout=900 > gpurun_out/g_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/g_pytest.log
tail -6 gpurun_out/g_pytest.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/g_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/g_smoke.log; tail -2 gpurun_out/g_smoke.log
timeout 900 python bench.py > gpurun_out/g_bench_n1.json 2> gpurun_out/g_bench_n1.err; echo "bench rc=$?"
tail -c 300 gpurun_out/g_bench_n1.err
timeout 600 ncu --set full --import-source on --clock-control none -k regex:rollout_tc2_kernel -c 1 -o gpurun_out/g_tc2 python tools/tc_check.py f16 2 > gpurun_out/g_ncu_tc2.log 2>&1
ncu -i gpurun_out/g_tc2.ncu-rep --page raw --csv > gpurun_out/g_tc2_raw.csv 2>/dev/null
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/g_launches_bench.csv python bench.py --steps 2 --warmup 3 > gpurun_out/g_ncu_bench.log 2>&1
ls -la gpurun_out/g_*
