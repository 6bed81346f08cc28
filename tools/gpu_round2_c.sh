#!/bin/bash
# GPU visit C (one GPU): single-load stem tiles - parity, bench, and an ncu capture of the fused stem + pool kernel
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 > gpurun_out/c_pytest.log; tail -3 gpurun_out/c_pytest.log
for tag in c1 c2; do
timeout 300 python bench.py --steps 60 --warmup 6 --no-cpu-baseline --sustained-seconds 3 --async-seconds 0 --dump-layers gpurun_out/${tag}_layers.json > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err
python - $tag <<'PY'
import json,sys
try:
    d=json.loads(open(f'gpurun_out/{sys.argv[1]}_bench.json').read().strip().splitlines()[-1])
    print(sys.argv[1], 'value', round(d['value']), 'e2e', round(d['e2e']['value']), 'sust', round(d['sustained']['value']), round(d['sustained']['e2e']), 'ms', round(d['ms_per_step'],4), {k:round(v,4) for k,v in d['layer_ms'].items()}, d['gpu_launches'])
except Exception as e: print(sys.argv[1], 'bench failed', e)
PY
done
AI4E_B200_NO_U8_STEM=1 timeout 300 python bench.py --steps 60 --warmup 6 --no-cpu-baseline --sustained-seconds 0 --async-seconds 0 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('K1 launched separately', round(d['value']), {k:round(v,4) for k,v in d['layer_ms'].items()})"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:stem_pool --launch-skip 4 -c 1 -f -o gpurun_out/c_stem_pool_u8 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --sustained-seconds 0 --async-seconds 0 > gpurun_out/c_ncu.txt 2>&1
ls -la gpurun_out/c_stem_pool_u8.ncu-rep | awk '{print $5}'
