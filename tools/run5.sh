#!/bin/bash
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_shard_engine.py -m gpu -q --timeout 180 -x 2>&1 | tail -3
run() {
  env "$@" timeout 300 python bench.py --steps 400 --warmup 40 --quick 2>/dev/null | python -c "
import json,sys
r=json.loads(sys.stdin.read().strip().splitlines()[-1]); p=r['phases']
print('$*: ms/step %.4f frac %.3f | gather %.2f plan %.2f update %.2f serial %.2f | parity %s' % (r['ms_per_step'], r['roofline']['step']['frac'], p['gather_us'], p['sort_plan_us'], p['reduce_update_us'], p['serial_step_us'], r['parity']['ok']))"
}
run X=1
run B200EMB_PDL=1
run B200EMB_REDUCE_CPS=2
run B200EMB_REDUCE_CPS=2 B200EMB_RING_KB=96
run B200EMB_REDUCE_CPS=2 B200EMB_GATHER_CPS=3
run B200EMB_REDUCE_CPS=2 B200EMB_GATHER_CPS=3 B200EMB_PDL=1
run B200EMB_REDUCE_CPS=2 B200EMB_RING_KB=96 B200EMB_GATHER_CPS=3 B200EMB_PDL=1
run B200EMB_REDUCE_CPS=2 B200EMB_GATHER_CPS=2
