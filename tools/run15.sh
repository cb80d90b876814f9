#!/bin/bash
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout 200 -x > gpurun_out/r2_pytest15.log 2>&1; echo "pytest rc=$?"
grep -E "passed|failed|Error" gpurun_out/r2_pytest15.log | tail -4 | cut -c1-300
run() {
  env "$@" timeout 300 python bench.py --steps 400 --warmup 40 --quick 2>/dev/null | python -c "
import json,sys
r=json.loads(sys.stdin.read().strip().splitlines()[-1]); p=r['phases']
print('$*: ms/step %.4f frac %.3f | gather %.2f plan %.2f update %.2f serial %.2f | parity %s' % (r['ms_per_step'], r['roofline']['step']['frac'], p['gather_us'], p['sort_plan_us'], p['reduce_update_us'], p['serial_step_us'], r['parity']['ok']))"
}
run X=1
run B200EMB_SORT_UNFUSED=1
run B200EMB_REDUCE_CPS=2
