#!/bin/bash
# 2-GPU visit, final tree: torchrun bench, one-process pool (closed loop) and native open loop
N=${1:-2}
mkdir -p gpurun_out
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 60 --warmup 5 --sustained-seconds 3 > gpurun_out/n${N}_bench.json 2> gpurun_out/n${N}_bench.err
python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/n${N}_bench.json').read().strip().splitlines()[-1]); print('torchrun N=${N}: value', round(d['value']), 'e2e', round(d['e2e']['value']), 'sustained', round(d['sustained']['value']), round(d['sustained']['e2e']), 'lat', d.get('latency_ms'))
except Exception as e: print('bench parse failed', e)
PY
timeout 200 python tools/async_bench.py --gpus $N --mode native --seconds 4 --clients $((N*2)) --out gpurun_out/n${N}_async.jsonl 2>&1 | tail -1 | cut -c1-300
PEAK=$(python - <<PY
import json
rows=[json.loads(l) for l in open('gpurun_out/n${N}_async.jsonl')]
print(int([r for r in rows if r['mode']=='native'][0]['tasks_per_s']))
PY
)
timeout 200 python tools/openloop_bench.py --gpus $N --peak $PEAK --load 0.5 0.8 0.9 --seconds 3 --feeders $((N*4)) --native --out gpurun_out/n${N}_openloop_native.jsonl 2>&1 | cut -c1-330
