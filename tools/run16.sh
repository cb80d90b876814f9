#!/bin/bash
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_fluid_ops.py tests/test_gpu_fullsize.py -m gpu -q --timeout 200 -x > gpurun_out/r2_pytest16.log 2>&1; echo "pytest rc=$?"
grep -E "passed|failed|Error" gpurun_out/r2_pytest16.log | tail -4 | cut -c1-300
rm -f gpurun_out/r2_phase16.jsonl
for ring in 1 0; do
  B200EMB_GATHER_RING=$ring timeout 300 python tools/phase_bench.py C2_ctr_sgd --tag ring$ring >> gpurun_out/r2_phase16.jsonl 2>/dev/null
  B200EMB_GATHER_RING=$ring timeout 300 python tools/phase_bench.py C3_adam --reps 40 --tag ring$ring >> gpurun_out/r2_phase16.jsonl 2>/dev/null
done
python - <<'PY'
import json
for l in open('gpurun_out/r2_phase16.jsonl'):
    r=json.loads(l); print(r['config'], r['tag'], 'gather %.2f (%.3f) plan %.1f update %.1f serial %.1f' % (r['gather_us'], r['gather_frac'], r['sort_plan_us'], r['update_us'], r['serial_step_us']))
PY
run() {
  env "$@" timeout 300 python bench.py --steps 400 --warmup 40 --quick 2>/dev/null | python -c "
import json,sys
r=json.loads(sys.stdin.read().strip().splitlines()[-1]); p=r['phases']
print('$*: ms/step %.4f frac %.3f | gather %.2f plan %.2f update %.2f serial %.2f | parity %s' % (r['ms_per_step'], r['roofline']['step']['frac'], p['gather_us'], p['sort_plan_us'], p['reduce_update_us'], p['serial_step_us'], r['parity']['ok']))"
}
run B200EMB_GATHER_RING=1
run B200EMB_GATHER_RING=0
