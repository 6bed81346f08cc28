#!/bin/bash
# GPU visit B (one GPU): the fused stem + max-pool kernel and the packed w3 stages - parity first, then the bench
# with and without each, the K1 stand-alone figures (events and ncu), the open-loop latency run
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "stem_pool or bottleneck_tail or maxpool or stem" 2>&1 | tail -8
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 > gpurun_out/b_pytest.log; tail -4 gpurun_out/b_pytest.log
run() { # tag, env...
  local tag=$1; shift
  env "$@" timeout 300 python bench.py --steps 60 --warmup 6 --no-cpu-baseline --sustained-seconds 3 --async-seconds 0 --dump-layers gpurun_out/b_layers_$tag.json > gpurun_out/b_bench_$tag.json 2> gpurun_out/b_bench_$tag.err
  python - $tag <<'PY'
import json,sys
try:
    d=json.loads(open(f'gpurun_out/b_bench_{sys.argv[1]}.json').read().strip().splitlines()[-1])
    print(sys.argv[1], 'value', round(d['value']), 'e2e', round(d['e2e']['value']), 'sust', round(d['sustained']['value']), round(d['sustained']['e2e']), 'ms', round(d['ms_per_step'],4), {k:round(v,4) for k,v in d['layer_ms'].items()}, d['gpu_launches'])
except Exception as e: print(sys.argv[1], 'bench failed', e)
PY
}
run fused A=1
run nostempool AI4E_B200_NO_STEM_POOL=1
run fused2 A=1
timeout 200 python tools/preprocess_bench.py --out gpurun_out/b_preprocess_events.json | cut -c1-200
timeout 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file gpurun_out/b_preprocess_ncu.csv python tools/preprocess_bench.py --once > /dev/null 2>&1
grep -c preprocess gpurun_out/b_preprocess_ncu.csv
PEAK=$(python -c "import json;print(int(json.loads(open('gpurun_out/b_bench_fused.json').read().strip().splitlines()[-1])['sustained']['e2e']))")
timeout 200 python tools/openloop_bench.py --gpus 1 --peak $PEAK --load 0.5 0.8 --seconds 4 --out gpurun_out/b_openloop.jsonl 2>&1 | tail -2 | cut -c1-330
