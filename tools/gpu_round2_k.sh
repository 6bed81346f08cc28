#!/bin/bash
# GPU visit K (one GPU): A/B of the MMA issue loop with descriptor low words kept in registers ("lean", ab/libai4e_lean.so)
# against the previous loop (ab/libai4e_base.so), each with the plain and the interleaved unit order of the fused tail
TAG=${1:-k}
mkdir -p gpurun_out
AI4E_B200_LIB=$PWD/ab/libai4e_lean.so timeout 600 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_engine.py -m gpu -x -q 2>&1 | tail -3 > gpurun_out/${TAG}_pytest_lean.log; tail -2 gpurun_out/${TAG}_pytest_lean.log
AI4E_B200_FUSED_INTERLEAVE=1 AI4E_B200_LIB=$PWD/ab/libai4e_lean.so timeout 300 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k bottleneck_tail 2>&1 | tail -2
for rep in 1 2; do
for v in base:0 lean:0 lean:1 lean:2; do
  lib=${v%%:*}; m=${v##*:}
  AI4E_B200_LIB=$PWD/ab/libai4e_$lib.so AI4E_B200_FUSED_INTERLEAVE=$m timeout 300 python bench.py --steps 100 --warmup 10 --no-cpu-baseline --sustained-seconds 0 --async-seconds 0 \
     --dump-layers gpurun_out/${TAG}_layers_${lib}_il$m.json 2>/dev/null | tail -1 > gpurun_out/${TAG}_bench_${lib}_il$m.json
  python - $TAG $lib $m <<'PY'
import json,sys
tag,lib,m=sys.argv[1:4]
try:
    d=json.loads(open(f'gpurun_out/{tag}_bench_{lib}_il{m}.json').read().strip().splitlines()[-1])
    L=json.load(open(f'gpurun_out/{tag}_layers_{lib}_il{m}.json'))['layers']
    t=lambda pre,keys: round(sum(r['ms'] for r in L if r['name'].startswith(pre) and any(k in r['name'] for k in keys))*1e3,1)
    print(lib,'interleave',m,'value',round(d['value']),'e2e',round(d['e2e']['value']),'parity',d['parity_checked'],'igemm_ms',round(d['layer_ms']['igemm_convs'],4),
          'l1 tails',t('layer1',('conv2','conv3')),'l2 tails',t('layer2',('conv2','conv3')),'l3',t('layer3',('conv','down')),'l4',t('layer4',('conv','down')))
except Exception as e: print('parse failed', lib, m, e)
PY
done
done
