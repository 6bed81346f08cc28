#!/bin/bash
# GPU visit L (one GPU): A/B of the hoisted next-stage wait in the MMA issue loops (ab/libai4e_hoist.so vs ab/libai4e_nohoist.so)
TAG=${1:-l}
mkdir -p gpurun_out
AI4E_B200_LIB=$PWD/ab/libai4e_hoist.so timeout 600 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_engine.py -m gpu -x -q 2>&1 | tail -3 > gpurun_out/${TAG}_pytest_hoist.log; tail -2 gpurun_out/${TAG}_pytest_hoist.log
for rep in 1 2; do
for lib in nohoist hoist; do
  AI4E_B200_LIB=$PWD/ab/libai4e_$lib.so timeout 300 python bench.py --steps 100 --warmup 10 --no-cpu-baseline --sustained-seconds 0 --async-seconds 0 \
     --dump-layers gpurun_out/${TAG}_layers_${lib}.json 2>/dev/null | tail -1 > gpurun_out/${TAG}_bench_${lib}.json
  python - $TAG $lib <<'PY'
import json,sys
tag,lib=sys.argv[1:3]
try:
    d=json.loads(open(f'gpurun_out/{tag}_bench_{lib}.json').read().strip().splitlines()[-1])
    L=json.load(open(f'gpurun_out/{tag}_layers_{lib}.json'))['layers']
    t=lambda pre,keys: round(sum(r['ms'] for r in L if r['name'].startswith(pre) and any(k in r['name'] for k in keys))*1e3,1)
    print(lib,'value',round(d['value']),'e2e',round(d['e2e']['value']),'parity',d['parity_checked'],'igemm_ms',round(d['layer_ms']['igemm_convs'],4),
          'l1 tails',t('layer1',('conv2','conv3')),'l2 tails',t('layer2',('conv2','conv3')),'l1 other',t('layer1',('conv1','down')),'l2 other',t('layer2',('conv1','down')),'l3',t('layer3',('conv','down')),'l4',t('layer4',('conv','down')))
except Exception as e: print('parse failed', lib, e)
PY
done
done
