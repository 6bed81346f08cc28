#!/bin/bash
# same-box A/B of several builds of libai4e_b200.so: tools/ab_bench.sh lib1.so lib2.so ...
for lib in "$@"; do
  for nc in 0; do
    if [ $nc = 1 ]; then export AI4E_B200_NO_CLUSTER=1; else unset AI4E_B200_NO_CLUSTER; fi
    AI4E_B200_LIB=$PWD/$lib timeout 100 python bench.py --steps 40 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.load(sys.stdin); print('$lib nocluster=$nc', round(d['value']), round(d['ms_per_step'],4), 'convs', round(d['layer_ms']['igemm_convs'],4))"
  done
done
