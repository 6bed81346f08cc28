#!/bin/bash
# tools/trace_shapes.sh lib.so  -> gpurun_out/trace_<name>.txt for a few ResNet-50 shapes (needs a -DAI4E_PIPE_TRACE build)
lib=$1
run() { name=$1; shift; AI4E_B200_LIB=$PWD/$lib AI4E_B200_PIPE_TRACE=$PWD/gpurun_out/trace_$name.txt timeout 60 python tools/conv_microbench.py "$@" 3 | tail -1; }
run l1ds   64 56 56 64 256 1 1 0 0 0
run l3c3   64 14 14 256 1024 1 1 0 1 1
run l1c1   64 56 56 256 64 1 1 0 1 0
