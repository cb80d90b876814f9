#!/bin/bash
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_shard_engine.py -m gpu -q --timeout 200 -k one_process > gpurun_out/r2_pytest_2gpu.log 2>&1; echo "pytest rc=$?"
grep -E "AssertionError|passed|failed|step [0-9]+ rank|PASSED|FAILED" gpurun_out/r2_pytest_2gpu.log | tail -8 | cut -c1-400
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 200 --warmup 20 --quick > gpurun_out/r2_sh2_quick.json 2> gpurun_out/r2_sh2_quick.err; echo "quick rc=$?"
grep -v "^W1107\|OMP_NUM\|^\*\*\*" gpurun_out/r2_sh2_quick.err | tail -6 | cut -c1-400
python - <<'PY'
import json
try:
    r=json.loads(open('gpurun_out/r2_sh2_quick.json').read().strip().splitlines()[-1])
    print('value %.3g ms/step %.4f graphs %s' % (r['value'], r['ms_per_step'], r['sweep'][0]['cuda_graphs']))
    print('phases', r['phases_us']); print('e2e', r['e2e']['value'], r['e2e']['ms_per_step']); print('parity', r['parity'])
except Exception as e:
    print('parse failed', e)
PY
