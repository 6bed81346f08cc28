#!/bin/bash
# ncu --set full of the two fused bottleneck tails on the final tree (leaner MMA loop, interleaved passes at Cmid 128)
mkdir -p gpurun_out
cap() {  # name, kernel regex, command...
  local name=$1 rx=$2; shift 2
  "$@" > gpurun_out/r02_v_${name}.plain.txt 2>&1 || { echo "$name: plain run failed"; return; }
  tail -1 gpurun_out/r02_v_${name}.plain.txt
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:$rx --launch-skip 3 -c 1 -f -o gpurun_out/r02_v_${name} "$@" > gpurun_out/r02_v_${name}.ncu.txt 2>&1
  ls -la gpurun_out/r02_v_${name}.ncu-rep 2>/dev/null | awk '{print $5, $9}'
}
cap l2tail bottleneck_tail python tools/fused_microbench.py 64 28 28 128 128 512 1 4
cap l1tail bottleneck_tail python tools/fused_microbench.py 64 56 56 64 64 256 1 4
