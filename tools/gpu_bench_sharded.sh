#!/bin/bash
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
N=${1:-2}
START=$(date +%s); timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 200 --warmup 20 > gpurun_out/r2_sh${N}_full.json 2> gpurun_out/r2_sh${N}_full.err; echo "full rc=$? wall=$(( $(date +%s) - START ))s"
grep -E "Error|error|FAILED" gpurun_out/r2_sh${N}_full.err | tail -5 | cut -c1-300
python - <<PY
import json
try:
    r=json.loads(open('gpurun_out/r2_sh${N}_full.json').read().strip().splitlines()[-1])
    print('value %.3g ms/step %.4f' % (r['value'], r['ms_per_step']), 'e2e %.3g' % r['e2e']['value'])
    for e in r['sweep']:
        print(e['ids_per_gpu_per_step'], 'ms %.4f value %.3g' % (e['ms_per_step'], e['value']), 'one_gpu', e.get('one_gpu_same_batch'), 'speedup', e.get('speedup_vs_one_gpu_same_batch'), 'hbm %.3f nvl %.3f' % (e['roofline']['hbm_frac'], e['roofline']['nvlink_frac']), 'graphs', e['cuda_graphs'])
    print('parity', {k:v for k,v in r['parity'].items() if k!='note'})
except Exception as e:
    print('parse failed', e)
PY
python - <<PY
import json
try:
    r=json.loads(open('gpurun_out/r2_sh${N}_full.json').read().strip().splitlines()[-1])
    for e in r.get('configs_extra', []):
        print(e['config'], 'ms %.3f value %.3g' % (e['ms_per_step'], e['value']), 'unique', e['unique_ids_per_gpu'], 'recv', e['rows_received_per_gpu'], 'hbm %.3f nvl %.3f' % (e['roofline']['hbm_frac'], e['roofline']['nvlink_frac']), 'graphs', e['cuda_graphs'])
        print('   parity', {k: v for k, v in e['parity'].items() if k != 'checked'})
        for k, v in e['phases_us'].items(): print('   %8.1f us  %s' % (v, k))
except Exception as e:
    print('parse failed', e)
PY
tail -5 gpurun_out/r2_sh${N}_full.err | cut -c1-300
