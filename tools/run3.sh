#!/bin/bash
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout 180 -x > gpurun_out/r2_pytest3.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest3.log
tail -8 gpurun_out/r2_pytest3.log | cut -c1-200
python tools/trace_reduce.py --launches 2 > gpurun_out/r2_trace3.log 2>&1; tail -22 gpurun_out/r2_trace3.log
