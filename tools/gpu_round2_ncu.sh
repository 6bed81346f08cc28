#!/bin/bash
# ncu evidence (one GPU): launch list of the bench command with DRAM bytes, and --set full captures of the
# kernels VERDICT names (layer3.1.conv1, layer4.0.conv1, a layer3 conv3 with residual, the layer1/2 tails)
mkdir -p gpurun_out
TAG=${1:-r02_a}
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --sustained-seconds 0 --async-seconds 0 > gpurun_out/${TAG}_ncu_cmd_plain.json 2>/dev/null || exit 1
AI4E_B200_NO_AUTOTUNE=1 timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 600 --csv \
  --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --sustained-seconds 0 --async-seconds 0 > gpurun_out/${TAG}_ncu.log 2>&1
tail -1 gpurun_out/${TAG}_launches.csv | cut -c1-160
cap() {  # name, kernel regex, command...
  local name=$1 rx=$2; shift 2
  "$@" > gpurun_out/${TAG}_${name}.plain.txt 2>&1 || { echo "$name: plain run failed"; return; }
  tail -1 gpurun_out/${TAG}_${name}.plain.txt
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:$rx --launch-skip 3 -c 1 -f -o gpurun_out/${TAG}_${name} "$@" > gpurun_out/${TAG}_${name}.ncu.txt 2>&1
  ls -la gpurun_out/${TAG}_${name}.ncu-rep 2>/dev/null | awk '{print $5, $9}'
}
cap l3c1 conv_igemm python tools/conv_microbench.py 64 14 14 1024 256 1 1 0 1 0 4
cap l4c1 conv_igemm python tools/conv_microbench.py 64 14 14 1024 512 1 1 0 1 0 4
cap l3c3 conv_igemm python tools/conv_microbench.py 64 14 14 256 1024 1 1 0 1 1 4
cap l3c2 conv_igemm python tools/conv_microbench.py 64 14 14 256 256 3 1 1 1 0 4
cap l2tail bottleneck_tail python tools/fused_microbench.py 64 28 28 128 128 512 1 4
cap l1tail bottleneck_tail python tools/fused_microbench.py 64 56 56 64 64 256 1 4
