#!/bin/bash
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python bench.py --steps 400 --warmup 40 > gpurun_out/r2_bench4.json 2> gpurun_out/r2_bench4.err; echo "bench rc=$?"
tail -3 gpurun_out/r2_bench4.err | cut -c1-300
python - <<'PY'
import json
try:
    r=json.loads(open('gpurun_out/r2_bench4.json').read().strip().splitlines()[-1])
    print('value %.3g ids/s  ms/step %.4f  step frac %.3f' % (r['value'], r['ms_per_step'], r['roofline']['step']['frac']))
    print('phases', {k:(round(v,2) if isinstance(v,float) else v) for k,v in r['phases'].items() if k!='note'})
    print('det_ms', r['det_ms_per_step'], 'e2e', r['e2e']['value'], 'e2e_opkernel', r['e2e_opkernel'])
    print('parity', r['parity'])
    for x in r['configs_extra']:
        print(x.get('config'), x.get('ms_per_step'), x.get('roofline',{}).get('step'), x.get('phases'), x.get('parity'), x.get('skipped'))
    print('clocks', r['clocks'])
except Exception as e:
    print('parse failed', e)
PY
for ch in 23; do
  B200EMB_CHUNK=$ch timeout 300 python bench.py --steps 400 --warmup 40 --quick 2>/dev/null | python -c "
import json,sys
r=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('chunk $ch: ms/step %.4f frac %.3f update %.2f' % (r['ms_per_step'], r['roofline']['step']['frac'], r['phases']['reduce_update_us']))"
done
