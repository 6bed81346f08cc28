#!/bin/bash
mkdir -p gpurun_out
lib=ab/libai4e_trace.so
run() { name=$1; shift; AI4E_B200_LIB=$PWD/$lib AI4E_B200_PIPE_TRACE=$PWD/gpurun_out/trace_$name.txt timeout 60 python tools/conv_microbench.py "$@" 3 | tail -1; }
run l1ds   64 56 56 64 256 1 1 0 0 0
run l3c3   64 14 14 256 1024 1 1 0 1 1
run l3c1   64 14 14 1024 256 1 1 0 1 0
run l3c2   64 14 14 256 256 3 1 1 1 0
run l2c1   64 28 28 512 128 1 1 0 1 0
python tools/trace_decode.py l1ds l3c3 l3c1 l3c2 l2c1 > gpurun_out/trace_decoded.txt 2>&1
