#!/bin/bash
# GPU visit A (one GPU): every GPU test, the bench line, the async path modes, land-cover at full size
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/a_pytest.log; tail -5 gpurun_out/a_pytest.log
timeout 400 python bench.py --dump-layers gpurun_out/a_layers.json > gpurun_out/a_bench.json 2> gpurun_out/a_bench.err; tail -c 600 gpurun_out/a_bench.err
python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/a_bench.json'))
    print('value', round(d['value']), 'e2e', round(d['e2e']['value']), 'ms', round(d['ms_per_step'],4), 'frac', round(d['roofline']['frac'],3), round(d['roofline']['frac_over_whole_step'],3))
    print('sustained', {k:(round(v) if isinstance(v,float) else v) for k,v in d.get('sustained',{}).items() if k in ('value','e2e')}, d.get('sustained',{}).get('clocks_value'))
    print('async', {k:(round(v.get('tasks_per_s',0)), v.get('p50_ms'), v.get('p99_ms')) if isinstance(v,dict) else v for k,v in d.get('async_path',{}).items()})
    print('cpu', d.get('cpu_baseline',{}).get('value'))
except Exception as e: print('bench parse failed', e)
PY
for m in blocked later native; do timeout 120 python tools/async_bench.py --gpus 1 --mode $m --seconds 5 --clients 64 --out gpurun_out/a_async.jsonl 2>&1 | tail -1 | cut -c1-400; done
timeout 120 python tools/async_bench.py --gpus 1 --mode native --seconds 5 --clients 4 --out gpurun_out/a_async.jsonl 2>&1 | tail -1 | cut -c1-300
timeout 300 python tools/landcover_bench.py --gpus 1 --seconds 4 --classify-rate 77000 --out gpurun_out/a_landcover.jsonl 2>&1 | tail -4 | cut -c1-330
