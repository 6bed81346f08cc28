#!/bin/bash
# quick GPU visit: selected kernel tests (arg 1 = -k expression), bench summary, ncu durations of the small kernels
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "$1" 2>&1 | tail -2
timeout 100 python bench.py --steps 40 --warmup 5 --no-cpu-baseline > gpurun_out/b.json 2>/dev/null
python - <<'PY'
import json; d=json.load(open("gpurun_out/b.json")); print(round(d["value"]), round(d["ms_per_step"],4), {k:round(v,4) for k,v in d["layer_ms"].items()})
PY
timeout 100 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"preprocess_lut|maxpool|avgpool|fc_kernel|softmax" -s 10 -c 5 python bench.py --steps 2 --warmup 3 --no-cpu-baseline 2>&1 | grep -E "gpu__time|^  [a-z].*\(" | paste - - | awk '{print $1, $2, $NF}'
