#!/bin/bash
# first GPU call of round 2: parity of the new reduce kernel, then phase timings
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest1.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest1.log
tail -5 gpurun_out/r2_pytest1.log
for kb in 100 110 72; do
  B200EMB_RING_KB=$kb timeout 300 python tools/phase_bench.py C2_ctr_sgd --tag ring$kb >> gpurun_out/r2_phase1.jsonl 2>> gpurun_out/r2_phase1.err
done
timeout 300 python tools/phase_bench.py C2_ctr_sgd --det --tag det >> gpurun_out/r2_phase1.jsonl 2>> gpurun_out/r2_phase1.err
for kb in 100 72; do
  B200EMB_RING_KB=$kb timeout 300 python tools/phase_bench.py C3_adam --reps 50 --tag ring$kb >> gpurun_out/r2_phase1.jsonl 2>> gpurun_out/r2_phase1.err
done
timeout 400 python tools/phase_bench.py C5_dup_heavy --reps 10 --tag default >> gpurun_out/r2_phase1.jsonl 2>> gpurun_out/r2_phase1.err
cat gpurun_out/r2_phase1.jsonl
tail -5 gpurun_out/r2_phase1.err
