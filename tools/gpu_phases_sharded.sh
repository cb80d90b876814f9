#!/bin/bash
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
N=${1:-8}
for ids in 1703936 106496; do
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 60 --warmup 10 --quick --ids-per-gpu $ids > gpurun_out/r2_sh${N}_ph_$ids.json 2> gpurun_out/r2_sh${N}_ph_$ids.err; echo "rc=$?"
python - <<PY
import json
try:
    r=json.loads(open('gpurun_out/r2_sh${N}_ph_$ids.json').read().strip().splitlines()[-1])
    print('$ids ids/GPU: ms/step %.4f value %.3g graphs %s recv_rows %s unique %s' % (r['ms_per_step'], r['value'], r['sweep'][0]['cuda_graphs'], r['config']['rows_received_per_gpu'], r['config']['unique_ids_per_gpu']))
    for k,v in r['phases_us'].items(): print('   %8.1f us  %s' % (v,k))
except Exception as e:
    print('parse failed', e)
PY
done
