#!/bin/bash
# one GPU visit: kernel tests, engine tests, bench, and the same bench with the env switch in $1 set (A/B)
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q 2>&1 | tail -2
timeout 400 python -m pytest tests/test_gpu_engine.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -3
for i in 1 2; do
AI4E_B200_TUNE_LOG=1 timeout 200 python bench.py --steps 40 --warmup 5 --no-cpu-baseline --dump-layers gpurun_out/layers.json 2>gpurun_out/tune.log > gpurun_out/bench_tuned.json
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_tuned.json')); print('default', round(d["value"]), round(d["ms_per_step"],4), 'e2e', round(d["e2e"]["value"]), {k:round(v,4) for k,v in d["layer_ms"].items()}, d["gpu_launches"])
PY
if [ -n "$1" ]; then
env $1=1 timeout 200 python bench.py --steps 40 --warmup 5 --no-cpu-baseline 2>/dev/null > gpurun_out/bench_alt.json
python - "$1" <<'PY'
import json,sys
d=json.load(open('gpurun_out/bench_alt.json')); print(sys.argv[1], round(d["value"]), round(d["ms_per_step"],4))
PY
fi
done
