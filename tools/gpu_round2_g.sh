#!/bin/bash
# GPU visit G (one GPU): smoke(), the full GPU suite, the land-cover bench with the K7 summary mode, a progressive-JPEG
# decode timing, then visit F's records (bench line, reference arm, per-layer table, ncu launch list)
TAG=${1:-g}
mkdir -p gpurun_out
timeout 300 python __graft_entry__.py --smoke > gpurun_out/${TAG}_smoke.log 2>&1; tail -2 gpurun_out/${TAG}_smoke.log
timeout 300 python tools/landcover_bench.py --gpus 1 --seconds 3 --classify-rate 80000 --out gpurun_out/${TAG}_landcover.jsonl > /dev/null 2> gpurun_out/${TAG}_landcover.err; tail -c 300 gpurun_out/${TAG}_landcover.err
python - $TAG <<'PY'
import json,sys
try:
    for l in open(f'gpurun_out/{sys.argv[1]}_landcover.jsonl'):
        d=json.loads(l); print(d['source_memory'], d['tiles_in_flight'], round(d['tiles_per_s'],1), d.get('frac_of_conv_bound'), d.get('summary_equals_aggregate_of_chip_results'), d['path'][:30])
except Exception as e: print('landcover parse failed', e)
PY
bash tools/gpu_round2_f.sh ${TAG}
