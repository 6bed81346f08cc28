#!/bin/bash
# same-box per-layer A/B of two builds: tools/ab_layers.sh libA.so libB.so  -> table of per-layer ms
i=0
for lib in "$@"; do
  AI4E_B200_LIB=$PWD/$lib timeout 100 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --dump-layers gpurun_out/abl_$i.json >/dev/null 2>&1
  i=$((i+1))
done
python - <<'PY'
import json
a=json.load(open('gpurun_out/abl_0.json'))['layers']; b=json.load(open('gpurun_out/abl_1.json'))['layers']
ta=tb=0
for x,y in zip(a,b):
    ta+=x['ms']; tb+=y['ms']
    print(f"{x['name']:28s} {str(x.get('in','')):18s} k{x.get('k','')} {x['ms']*1e3:7.1f} {y['ms']*1e3:7.1f} {(y['ms']/x['ms']-1)*100:+6.1f}%")
print('total', round(ta*1e3,1), round(tb*1e3,1))
PY
