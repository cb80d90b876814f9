#!/bin/bash
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q --timeout 180 -k "det or exact or parity or golden or fluid" > gpurun_out/r2_pytest18.log 2>&1; echo "pytest rc=$?"
grep -E "passed|failed|Error" gpurun_out/r2_pytest18.log | tail -8 | cut -c1-300
for c in C2_ctr_sgd C3_adam C5_dup_heavy; do timeout 200 python tools/phase_bench.py $c --reps 100 --det 2>&1 | tail -1 | cut -c1-900; done
