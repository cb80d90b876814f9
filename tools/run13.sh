#!/bin/bash
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout 200 -x > gpurun_out/r2_pytest13.log 2>&1; echo "pytest rc=$?"
grep -E "passed|failed|Error" gpurun_out/r2_pytest13.log | tail -4 | cut -c1-300
python tools/shard_phase1.py 2>&1 | tail -16
