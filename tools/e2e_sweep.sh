#!/bin/bash
for n in 2 3 4; do
AI4E_BENCH_E2E_INFLIGHT=$n timeout 200 python bench.py --steps 80 --warmup 6 --no-cpu-baseline 2>/dev/null > gpurun_out/e2e_$n.json
python - $n <<'PY'
import json,sys
d=json.load(open('gpurun_out/e2e_%s.json'%sys.argv[1])); print('inflight',sys.argv[1], 'value', round(d["value"]), 'e2e', round(d["e2e"]["value"]), d["latency_ms"])
PY
done
