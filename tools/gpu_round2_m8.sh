#!/bin/bash
# 8-GPU visit, final tree: one Pool over N GPUs in one process (native dispatch, open loop with enough feeder threads,
# land-cover with the K7 summary, micro-batch sweep) and the torchrun bench
N=${1:-8}
mkdir -p gpurun_out
nproc > gpurun_out/n${N}_nproc.txt
timeout 200 python tools/async_bench.py --gpus $N --mode native --seconds 5 --clients $((N*2)) --out gpurun_out/n${N}_async.jsonl 2>&1 | tail -1 | cut -c1-300
PEAK=$(python - <<PY
import json
rows=[json.loads(l) for l in open('gpurun_out/n${N}_async.jsonl')]
print(int([r for r in rows if r['mode']=='native'][0]['tasks_per_s']))
PY
)
timeout 200 python tools/openloop_bench.py --gpus $N --peak $PEAK --load 0.5 0.8 --seconds 4 --feeders $((N*3)) --out gpurun_out/n${N}_openloop.jsonl 2>&1 | tail -2 | cut -c1-420
timeout 300 python tools/landcover_bench.py --gpus $N --seconds 3 --classify-rate 79000 --out gpurun_out/n${N}_landcover.jsonl 2>&1 | tail -5 | cut -c1-260
timeout 400 python tools/extra_benches.py gpurun_out/n${N}_sweep.json --gpus $N --batches 1,8,64,512 > /dev/null 2> gpurun_out/n${N}_sweep.err
python - <<PY
import json
try:
    for r in json.load(open('gpurun_out/n${N}_sweep.json'))['batch_sweep']:
        print('sweep N=${N}', r.get('batch'), round(r.get('images_per_s',0)), round(r.get('e2e_images_per_s',0)), r.get('latency_ms',{}).get('p50'), r.get('latency_ms',{}).get('p99'), r.get('error','')[:100])
except Exception as e: print('sweep failed', e)
PY
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 60 --warmup 5 --sustained-seconds 3 > gpurun_out/n${N}_bench.json 2> gpurun_out/n${N}_bench.err
python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/n${N}_bench.json').read().strip().splitlines()[-1]); print('torchrun N=${N}: value', round(d['value']), 'e2e', round(d['e2e']['value']), 'sustained', round(d['sustained']['value']), round(d['sustained']['e2e']), 'lat', d.get('latency_ms'))
except Exception as e: print('bench parse failed', e)
PY
