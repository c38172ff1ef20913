#!/bin/bash
# correctness of the new build first (bit-equality of the launch variants and against the one-tile kernel), then a
# same-box A/B of the committed build (git worktree ab_old/) against the working tree
mkdir -p gpurun_out
timeout 240 python tools/tc2_check.py > gpurun_out/ab2_check.log 2>&1; rc=$?
grep -v "^cfg" gpurun_out/ab2_check.log | tail -7
if [ $rc -ne 0 ] && [ -z "$AB_FORCE" ]; then echo "tc2_check rc=$rc: skipping the A/B"; exit 1; fi
[ -n "$AB_FORCE" ] || timeout 200 python -m pytest tests/test_gpu_mpc.py -x -q 2>&1 | tail -3
show='
import sys, json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print(sys.argv[1], "value %.4g ms %.4f e2e %.4g kernel_ms %.4f sus %.4g" % (d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"]["kernel_ms"], d["sustained"]["candidate_steps_per_s"]),
      "verified %d/%d fallback %d margin,eta %s err_all %.3g err_sel %.3g" % (d["verify"]["verified1"], d["verify"]["plans"], d["verify"]["fallback2"], d["verify"]["last_plan_margin_eta"], d["parity"]["cost_rel_err_all_max"], d["parity"]["cost_rel_err_selected_max"]))'
(cd ab_old && timeout 200 python tools/tc2_sum.py 2>&1 | grep '^sum' > ../gpurun_out/sum_old.txt); timeout 200 python tools/tc2_sum.py 2>&1 | grep '^sum' > gpurun_out/sum_new.txt
if cmp -s gpurun_out/sum_old.txt gpurun_out/sum_new.txt; then echo "checksums: old == new on $(wc -l < gpurun_out/sum_new.txt) shapes"; else echo 'checksums DIFFER'; diff gpurun_out/sum_old.txt gpurun_out/sum_new.txt; fi
for rep in 1 2; do
(cd ab_old && timeout 300 python bench.py --steps 40 --warmup 5 2>/dev/null | python -c "$show" old_lib)
(timeout 300 python bench.py --steps 40 --warmup 5 2>/dev/null | python -c "$show" new_lib)
done
