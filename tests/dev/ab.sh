#!/bin/bash
# Development aid: A/B of several builds of the library inside ONE gpurun call (boxes differ by a few percent).
# usage: tests/dev/ab.sh [rounds]   (variants: tiny-cuda-nn_b200/lib/variants/*.so, plus the default library as "default")
for r in $(seq 1 ${1:-3}); do
  for lib in default $PWD/tiny-cuda-nn_b200/lib/variants/*.so; do
    if [ $lib = default ]; then unset TCNN_B200_LIB; else export TCNN_B200_LIB=$lib; fi
    python bench.py --steps 1000 --warmup 30 --no-cpu-baseline --e2e-steps 20 2>/dev/null | python -c "
import json,sys
d=[json.loads(l) for l in sys.stdin if l.startswith('{')][-1]
print('$(basename $lib)', round(d['ms_per_step'],4), {k:round(v,4) for k,v in d['roofline']['kernel_ms_per_step_serial'].items()})"
  done
done
