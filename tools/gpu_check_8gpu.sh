#!/bin/bash
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_shard_engine.py -m gpu -q --timeout 300 -k "one_process and 8" > gpurun_out/r2_pytest_8gpu.log 2>&1; echo "pytest rc=$?"
grep -E "passed|failed|step [0-9]+ rank" gpurun_out/r2_pytest_8gpu.log | tail -4 | cut -c1-600
bash tools/gpu_bench_sharded.sh 8
python - <<'PY'
import json
r=json.loads(open('gpurun_out/r2_sh8_full.json').read().strip().splitlines()[-1])
for e in r['sweep']:
    print(e['ids_per_gpu_per_step']); [print('   %8.1f us  %s' % (v,k)) for k,v in e['phases_us'].items()]
PY
