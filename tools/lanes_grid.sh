#!/bin/bash
for cfg in "2 132 3" "2 132 4" "3 100 4" "3 116 4" "3 88 4" "3 100 3"; do
set -- $cfg
AI4E_B200_NO_AUTOTUNE=1 AI4E_B200_LANES=$1 AI4E_B200_GRID_SMS=$2 AI4E_BENCH_E2E_INFLIGHT=$3 timeout 200 python bench.py --steps 80 --warmup 6 --no-cpu-baseline 2>/dev/null > gpurun_out/lg.json
python - $1 $2 $3 <<'PY'
import json,sys
d=json.load(open('gpurun_out/lg.json')); print('lanes',sys.argv[1],'grid',sys.argv[2],'e2e inflight',sys.argv[3], 'value', round(d["value"]), round(d["ms_per_step"],4), 'e2e', round(d["e2e"]["value"]), {k:round(v,2) for k,v in d["latency_ms"].items() if k!='of'})
PY
done
