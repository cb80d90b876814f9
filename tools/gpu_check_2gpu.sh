#!/bin/bash
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_shard_engine.py tests/test_gpu_sharded.py tests/test_fluid_ops.py -m gpu -q --timeout 200 > gpurun_out/r2_pytest_2gpu.log 2>&1; echo "pytest rc=$?"
grep -E "passed|failed|Error|step [0-9]+ rank" gpurun_out/r2_pytest_2gpu.log | tail -6 | cut -c1-400
bash tools/gpu_bench_sharded.sh 2
