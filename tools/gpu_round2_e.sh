#!/bin/bash
# GPU visit E (one GPU): balanced run lengths in the stem + pool kernel (parity + bench), then the micro-batch sweep 1..512
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 > gpurun_out/e_pytest.log; tail -3 gpurun_out/e_pytest.log
timeout 300 python bench.py --steps 60 --warmup 6 --no-cpu-baseline --sustained-seconds 3 --async-seconds 0 --dump-layers gpurun_out/e_layers.json > gpurun_out/e_bench.json 2> gpurun_out/e_bench.err
python - <<'PY'
import json
try:
    d=json.loads(open('gpurun_out/e_bench.json').read().strip().splitlines()[-1])
    print('e', 'value', round(d['value']), 'e2e', round(d['e2e']['value']), 'sust', round(d['sustained']['value']), round(d['sustained']['e2e']), 'ms', round(d['ms_per_step'],4), {k:round(v,4) for k,v in d['layer_ms'].items()}, d['gpu_launches'])
except Exception as e: print('bench failed', e)
PY
timeout 900 python tools/extra_benches.py gpurun_out/e_sweep.json > /dev/null 2> gpurun_out/e_sweep.err
python - <<'PY'
import json
try:
    for r in json.load(open('gpurun_out/e_sweep.json'))['batch_sweep']:
        print(r.get('batch'), round(r.get('images_per_s',0)), round(r.get('ms_per_step',0),4), round(r.get('e2e_images_per_s',0)), r.get('latency_ms'), r.get('error','')[:100])
except Exception as e: print('sweep failed', e)
PY
