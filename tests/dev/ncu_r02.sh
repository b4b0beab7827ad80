#!/bin/bash
# Development aid (GPU box): the ncu evidence of a round.  usage: tests/dev/ncu_r02.sh <outdir>
#   1. launch list (gpu__time_duration of every kernel of a short bench run) -> launches.csv
#   2. one --set full capture of each hot kernel (after warm-up)              -> full.ncu-rep + full_raw.csv
# Numbers printed by a run under ncu are never bench values.
out=${1:-gpurun_out/ncu}; mkdir -p $out
python bench.py --steps 20 --warmup 5 --no-cpu-baseline --e2e-steps 5 > $out/plain.json 2> $out/plain.err || exit 1
TCNN_B200_GRAPH=0 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/launches.csv \
  python bench.py --steps 20 --warmup 5 --no-cpu-baseline --e2e-steps 5 > $out/ncu_launches.log 2>&1
TCNN_B200_GRAPH=0 ncu --set full --clock-control none --import-source on -k regex:"mlp_kernel|grid_forward_kernel|adam_kernel|mlp_finalize_adam_kernel" -s 40 -c 4 -f -o $out/full \
  python bench.py --steps 20 --warmup 5 --no-cpu-baseline --e2e-steps 5 > $out/ncu_full.log 2>&1
ncu -i $out/full.ncu-rep --page raw --csv > $out/full_raw.csv 2> $out/full_raw.err
ls -la $out
