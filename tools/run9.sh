#!/bin/bash
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout 180 > gpurun_out/r2_pytest9.log 2>&1; echo "pytest rc=$?"
grep -E "passed|failed|Error" gpurun_out/r2_pytest9.log | tail -8 | cut -c1-300
timeout 300 python bench.py --steps 200 --warmup 20 --quick 2>gpurun_out/r2_bench9.err | python -c "
import json,sys
r=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('ms/step %.4f' % r['ms_per_step'], 'e2e_opkernel', json.dumps(r['e2e_opkernel'])[:600])"
tail -3 gpurun_out/r2_bench9.err | cut -c1-300
