#!/bin/bash
# GPU visit D (one GPU): 16-warp epilogue of the stem + pool kernel - parity, bench, e2e with 4 batches in flight
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 > gpurun_out/g_pytest.log; tail -3 gpurun_out/d_pytest.log
run() { local tag=$1; shift
env "$@" timeout 300 python bench.py --steps 60 --warmup 6 --no-cpu-baseline --sustained-seconds 3 --async-seconds 0 --dump-layers gpurun_out/${tag}_layers.json > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err
python - $tag <<'PY'
import json,sys
try:
    d=json.loads(open(f'gpurun_out/{sys.argv[1]}_bench.json').read().strip().splitlines()[-1])
    print(sys.argv[1], 'value', round(d['value']), 'e2e', round(d['e2e']['value']), 'sust', round(d['sustained']['value']), round(d['sustained']['e2e']), 'ms', round(d['ms_per_step'],4), {k:round(v,4) for k,v in d['layer_ms'].items()}, d['gpu_launches'])
except Exception as e: print(sys.argv[1], 'bench failed', e)
PY
}
run g1 A=1
run g2 A=1
for s in "64 56 56 64 256 1 1 0 0 0" "64 14 14 256 1024 1 1 0 1 1" "64 14 14 1024 256 1 1 0 1 0" "64 14 14 256 256 3 1 1 1 0" "64 28 28 512 128 1 1 0 1 0" "64 56 56 256 512 1 2 0 0 0"; do python tools/conv_microbench.py $s 10 | tail -1; done
