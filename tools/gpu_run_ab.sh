#!/bin/bash
mkdir -p gpurun_out
show='
import sys, json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print(sys.argv[1], "value %.4g ms %.4f e2e %.4g kernel_ms %.4f sus %.4g" % (d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"]["kernel_ms"], d["sustained"]["candidate_steps_per_s"]))'
for rep in 1 2; do
(cd ab_old && timeout 300 python bench.py --steps 40 --warmup 5 2>/dev/null | python -c "$show" old_lib)
(timeout 300 python bench.py --steps 40 --warmup 5 2>/dev/null | python -c "$show" new_lib)
done
timeout 240 python tools/tc2_check.py 2>&1 | grep -v "^cfg" | tail -7
