#!/bin/bash
# stage-count experiment: the fused tail at 2 / 3 / all stages (does time scale with bytes in flight?)
for lib in ab/libai4e_s2.so ab/libai4e_s3.so aiforearth-api-platform_b200/ai4e_b200/libai4e_b200.so; do
  echo "== $lib"
  AI4E_B200_LIB=$PWD/$lib python tools/fused_microbench.py 64 28 28 128 128 512 1 10
  AI4E_B200_LIB=$PWD/$lib python tools/fused_microbench.py 64 56 56 64 64 256 1 10
done
