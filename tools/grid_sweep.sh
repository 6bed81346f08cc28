#!/bin/bash
for n in 148 140 136 132 128 124 148 132; do
AI4E_B200_NO_AUTOTUNE=1 AI4E_B200_GRID_SMS=$n timeout 200 python bench.py --steps 60 --warmup 6 --no-cpu-baseline 2>/dev/null > gpurun_out/grid_$n.json
python - $n <<'PY'
import json,sys
d=json.load(open('gpurun_out/grid_%s.json'%sys.argv[1])); print('grid sms',sys.argv[1], 'value', round(d["value"]), round(d["ms_per_step"],4), 'e2e', round(d["e2e"]["value"]))
PY
done
