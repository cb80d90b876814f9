#!/bin/bash
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout 180 > gpurun_out/r2_pytest2.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest2.log
tail -15 gpurun_out/r2_pytest2.log
python tools/trace_reduce.py --launches 2 > gpurun_out/r2_trace2.log 2>&1; tail -22 gpurun_out/r2_trace2.log
rm -f gpurun_out/r2_phase2.jsonl
for ch in 0 10 20; do
  B200EMB_CHUNK=$ch timeout 300 python tools/phase_bench.py C2_ctr_sgd --tag chunk$ch >> gpurun_out/r2_phase2.jsonl 2>> gpurun_out/r2_phase2.err
done
timeout 300 python tools/phase_bench.py C3_adam --reps 50 --tag default >> gpurun_out/r2_phase2.jsonl 2>> gpurun_out/r2_phase2.err
timeout 400 python tools/phase_bench.py C5_dup_heavy --reps 10 --tag default >> gpurun_out/r2_phase2.jsonl 2>> gpurun_out/r2_phase2.err
python - <<'PY'
import json
for l in open('gpurun_out/r2_phase2.jsonl'):
    r=json.loads(l); print(r['config'], r['tag'], 'gather %.1f plan %.1f update %.1f (%.3f) serial %.1f' % (r['gather_us'], r['sort_plan_us'], r['update_us'], r['update_frac'], r['serial_step_us']))
PY
tail -5 gpurun_out/r2_phase2.err
