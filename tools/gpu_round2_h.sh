#!/bin/bash
# GPU visit H (one GPU): the interleaved unit order of the fused bottleneck tail (AI4E_B200_FUSED_INTERLEAVE = 0 / 1 / 2):
# kernel and engine parity with the order on, then value / per-layer times of each mode from the same box
TAG=${1:-h}
mkdir -p gpurun_out
for m in 1 2; do
  AI4E_B200_FUSED_INTERLEAVE=$m timeout 300 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_engine.py -m gpu -x -q -k "bottleneck_tail or forward_host_matches_golden or landcover_4band or stages_match or two_lanes" 2>&1 | tail -3 > gpurun_out/${TAG}_pytest_il$m.log
  tail -2 gpurun_out/${TAG}_pytest_il$m.log
done
for m in 0 1 2 0 1 2; do
  AI4E_B200_FUSED_INTERLEAVE=$m timeout 300 python bench.py --steps 100 --warmup 10 --no-cpu-baseline --sustained-seconds 0 --async-seconds 0 \
     --dump-layers gpurun_out/${TAG}_layers_il$m.json 2>/dev/null | tail -1 > gpurun_out/${TAG}_bench_il$m.json
  python - $TAG $m <<'PY'
import json,sys
tag,m=sys.argv[1:3]
try:
    d=json.loads(open(f'gpurun_out/{tag}_bench_il{m}.json').read().strip().splitlines()[-1])
    L=json.load(open(f'gpurun_out/{tag}_layers_il{m}.json'))['layers']
    t=lambda pre: round(sum(r['ms'] for r in L if r['name'].startswith(pre) and ('conv2' in r['name'] or 'conv3' in r['name']))*1e3,1)
    print('interleave',m,'value',round(d['value']),'e2e',round(d['e2e']['value']),'parity',d['parity_checked'],'igemm_ms',round(d['layer_ms']['igemm_convs'],4),'l1 tails',t('layer1'),'l2 tails',t('layer2'))
except Exception as e: print('parse failed', m, e)
PY
done
