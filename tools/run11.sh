#!/bin/bash
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_fluid_ops.py -m gpu -q --timeout 180 > gpurun_out/r2_pytest11.log 2>&1; echo "pytest rc=$?"
grep -E "passed|failed|Error|assert" gpurun_out/r2_pytest11.log | tail -12 | cut -c1-300
