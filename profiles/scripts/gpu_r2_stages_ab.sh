#!/bin/bash
# A/B of stage-kernel build variants: ncu durations of the render-chunk stage kernels per library
set -u
mkdir -p gpurun_out
M=gpu__time_duration.sum
for v in "" _exp0_pdf6 _exp0_pdf10 _exp0_ur1 _exp0_ur6; do
  export NERF_B200_LIB=$PWD/nerf-replication_b200/libnerf_b200$v.so
  timeout 600 ncu --metrics $M --clock-control none -k regex:"sample_pdf|composite" -c 16 --csv --log-file gpurun_out/stg_ab.csv \
    python bench.py --mode render --steps 1 --warmup 1 --no-gpu-baseline --no-cpu-baseline > /dev/null 2>&1
  echo "== variant '$v'"; python profiles/summarize.py launches gpurun_out/stg_ab.csv | tail -3
done
