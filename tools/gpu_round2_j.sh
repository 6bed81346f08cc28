#!/bin/bash
# GPU visit J: pipeline traces of the fused bottleneck tail (CTA 0) at the layer1 / layer2 shapes, plain and interleaved order
mkdir -p gpurun_out
lib=${TRACE_LIB:-ab/libai4e_trace1.so}
run() { name=$1; il=$2; shift 2; AI4E_B200_FUSED_INTERLEAVE=$il AI4E_B200_LIB=$PWD/$lib AI4E_B200_PIPE_TRACE=$PWD/gpurun_out/ftrace_${name}_il$il.txt timeout 60 python tools/fused_microbench.py "$@" 3 | tail -1; }
for il in 0 1; do
  run l2tail $il 64 28 28 128 128 512 1
  run l1tail $il 64 56 56 64 64 256 1
done
for f in gpurun_out/ftrace_*.txt; do python tools/trace_decode_fused.py $f > ${f%.txt}_decoded.txt 2>&1; done
tail -25 gpurun_out/ftrace_l2tail_il0_decoded.txt
