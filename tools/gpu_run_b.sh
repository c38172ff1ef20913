#!/bin/bash
# round-2 GPU call B: fixed probe, whole GPU test-suite (no -x), ndiv timing + launch list, default bench
mkdir -p gpurun_out
timeout 120 ./tools/tmem_probe > gpurun_out/b_tmem_probe.txt 2>&1; echo "probe rc=$?" >> gpurun_out/b_tmem_probe.txt
timeout 2400 python -m pytest tests -m gpu -q --timeout=900 > gpurun_out/b_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/b_pytest.log
tail -15 gpurun_out/b_pytest.log
timeout 300 python tools/ndiv_time.py cfg3 > gpurun_out/b_ndiv_time.txt 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/b_launches_ndiv_cfg3.csv python tools/ndiv_time.py cfg3 > /dev/null 2>&1
timeout 900 python bench.py > gpurun_out/b_bench_n1.json 2> gpurun_out/b_bench_n1.err; echo "bench rc=$?"
tail -c 400 gpurun_out/b_bench_n1.err
