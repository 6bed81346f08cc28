#!/bin/bash
# one GPU visit: kernel tests, bench with the autotuned tilings (+ log), bench with the cost model only, engine tests
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q 2>&1 | tail -3
AI4E_B200_TUNE_LOG=1 timeout 200 python bench.py --steps 40 --warmup 5 --no-cpu-baseline 2>gpurun_out/tune.log > gpurun_out/bench_tuned.json
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_tuned.json')); print('tuned', round(d["value"]), round(d["ms_per_step"],4), 'e2e', round(d["e2e"]["value"]), d["layer_ms"])
PY
AI4E_B200_NO_AUTOTUNE=1 timeout 200 python bench.py --steps 40 --warmup 5 --no-cpu-baseline 2>/dev/null > gpurun_out/bench_model.json
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_model.json')); print('model only', round(d["value"]), round(d["ms_per_step"],4))
PY
timeout 400 python -m pytest tests/test_gpu_engine.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -3
