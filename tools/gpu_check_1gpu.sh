#!/bin/bash
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout 180 > gpurun_out/r2_pytest17.log 2>&1; echo "pytest rc=$?"
grep -E "passed|failed|Error" gpurun_out/r2_pytest17.log | tail -8 | cut -c1-300
timeout 400 python bench.py --steps 200 --warmup 20 > gpurun_out/r2_bench17.json 2>gpurun_out/r2_bench17.err; echo "bench rc=$?"
python -c "
import json
r=json.loads(open('gpurun_out/r2_bench17.json').read().strip().splitlines()[-1])
print('ms/step %.4f' % r['ms_per_step'], 'det', r.get('det_ms_per_step'), 'value %.3g' % r['value'])
print(json.dumps(r.get('parity'))[:300])
for c in r.get('configs_extra', []): print(json.dumps(c)[:400])
"
tail -3 gpurun_out/r2_bench17.err | cut -c1-300
