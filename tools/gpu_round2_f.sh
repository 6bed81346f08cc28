#!/bin/bash
# GPU visit F (one GPU): full GPU test suite, then the complete default bench line (value, e2e, sustained, async_path,
# cpu_baseline), the reference arm, the per-layer table, and the ncu launch list of the bench command
TAG=${1:-f}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 > gpurun_out/${TAG}_pytest.log; tail -3 gpurun_out/${TAG}_pytest.log
timeout 600 python bench.py --dump-layers gpurun_out/${TAG}_layers.json > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; tail -c 300 gpurun_out/${TAG}_bench.err
python - $TAG <<'PY'
import json,sys
try:
    d=json.loads(open(f'gpurun_out/{sys.argv[1]}_bench.json').read().strip().splitlines()[-1])
    print('value', round(d['value']), 'e2e', round(d['e2e']['value']), 'ms', round(d['ms_per_step'],4), 'frac', round(d['roofline']['frac'],3), round(d['roofline']['frac_over_whole_step'],3), round(d['roofline']['frac_over_whole_step_vs_burst'],3))
    print('sustained', round(d['sustained']['value']), round(d['sustained']['e2e']), d['sustained']['clocks_value'])
    print('layers', {k:round(v,4) for k,v in d['layer_ms'].items()}, 'launches', d['gpu_launches'], 'parity', d['parity_checked'])
    print('async', {k:(round(v.get('images_per_s',0)), v.get('p50_ms'), v.get('p99_ms')) if isinstance(v,dict) else v for k,v in d.get('async_path',{}).items()})
    print('cpu', d.get('cpu_baseline',{}).get('value'), 'pre', d['roofline_preprocess']['frac'])
except Exception as e: print('bench parse failed', e)
PY
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_reference.json 2>/dev/null; tail -c 250 gpurun_out/${TAG}_reference.json; echo
AI4E_B200_NO_AUTOTUNE=1 timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 600 --csv \
  --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --sustained-seconds 0 --async-seconds 0 > gpurun_out/${TAG}_ncu.log 2>&1
tail -1 gpurun_out/${TAG}_launches.csv | cut -c1-120
