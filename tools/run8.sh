#!/bin/bash
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
nvidia-smi -L | wc -l
timeout 500 python -m pytest tests/test_gpu_shard_engine.py -m gpu -q --timeout 300 -k one_process > gpurun_out/r2_pytest_8gpu.log 2>&1; echo "pytest rc=$?"
grep -E "passed|failed|step [0-9]+ rank" gpurun_out/r2_pytest_8gpu.log | tail -4 | cut -c1-600
bash tools/run7.sh 8
bash tools/run7.sh 4
