#!/bin/bash
# round-2 GPU call A: probes, GPU test-suite, default bench (outputs under gpurun_out/)
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv > gpurun_out/a_gpu.txt 2>&1
timeout 120 ./tools/tmem_probe > gpurun_out/a_tmem_probe.txt 2>&1; echo "probe rc=$?" >> gpurun_out/a_tmem_probe.txt
timeout 1500 python -m pytest tests -m gpu -q -x --timeout=600 > gpurun_out/a_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/a_pytest.log
tail -5 gpurun_out/a_pytest.log
timeout 900 python bench.py > gpurun_out/a_bench_n1.json 2> gpurun_out/a_bench_n1.err; echo "bench rc=$?"
tail -c 600 gpurun_out/a_bench_n1.err
