#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_mpc.py -x -q -k "rollout_matches_oracle or rejects_unsupported or embedding_256 or module_semantics or plan_golden" 2>&1 | tail -5
timeout 200 python tools/e256_time.py > gpurun_out/e256_time.json 2> gpurun_out/e256_time.err; echo "e256_time rc=$?"; cat gpurun_out/e256_time.json; tail -3 gpurun_out/e256_time.err
