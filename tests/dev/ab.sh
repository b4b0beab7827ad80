#!/bin/bash
# Development aid: A/B of two builds of the library inside ONE gpurun call (boxes differ by a few percent).
# usage: tests/dev/ab.sh [rounds]   (B = tiny-cuda-nn_b200/lib/libtcnn_b200_alt.so)
for r in $(seq 1 ${1:-3}); do
  for v in A B; do
    if [ $v = B ]; then export TCNN_B200_LIB=$PWD/tiny-cuda-nn_b200/lib/libtcnn_b200_alt.so; else unset TCNN_B200_LIB; fi
    python bench.py --steps 1000 --warmup 30 --no-cpu-baseline --e2e-steps 20 2>/dev/null | python -c "
import json,sys
d=[json.loads(l) for l in sys.stdin if l.startswith('{')][-1]
print('$v', round(d['ms_per_step'],4), {k:round(v,4) for k,v in d['roofline']['kernel_ms_per_step_serial'].items()})"
  done
done
