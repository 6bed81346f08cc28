#!/bin/bash
for n in 1 2 3 4; do
AI4E_B200_LANES=$n timeout 200 python bench.py --steps 60 --warmup 6 --no-cpu-baseline 2>/dev/null > gpurun_out/lanes_$n.json
python - $n <<'PY'
import json,sys
d=json.load(open('gpurun_out/lanes_%s.json'%sys.argv[1])); print('lanes',sys.argv[1], round(d["value"]), round(d["ms_per_step"],4), 'e2e', round(d["e2e"]["value"]), d["latency_ms"])
PY
done
