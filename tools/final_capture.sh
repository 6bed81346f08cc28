#!/bin/bash
# end-of-round evidence: full bench (with cpu_baseline), reference arm, per-layer table, ncu launch list with DRAM bytes
mkdir -p gpurun_out
timeout 300 python bench.py --steps 60 --warmup 6 --dump-layers gpurun_out/r01_k_layers.json > gpurun_out/r01_k_bench.json 2>gpurun_out/bench.err
tail -c 400 gpurun_out/r01_k_bench.json; echo
timeout 200 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r01_k_reference.json 2>/dev/null
tail -c 300 gpurun_out/r01_k_reference.json; echo
AI4E_B200_NO_AUTOTUNE=1 timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 300 --csv --log-file gpurun_out/r01_k_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_k.log 2>&1
tail -2 gpurun_out/r01_k_launches.csv | cut -c1-200
