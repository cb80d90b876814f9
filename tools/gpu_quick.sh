#!/bin/bash
# quick look after a kernel change: parity tests of the reduce paths, then phase times of C2 / C3
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py -m gpu -q --timeout 180 -x > gpurun_out/quick_pytest.log 2>&1; echo "pytest rc=$?"
grep -E "passed|failed|Error" gpurun_out/quick_pytest.log | tail -4 | cut -c1-300
for c in C2_ctr_sgd C3_adam; do timeout 200 python tools/phase_bench.py $c --reps 200 2>&1 | tail -1 | python -c "
import json,sys
r=json.loads(sys.stdin.read()); print(r['config'], 'gather %.1f plan %.1f update %.2f (%.3f) serial %.1f' % (r['gather_us'], r['sort_plan_us'], r['update_us'], r['update_frac'], r['serial_step_us']))"; done
