#!/bin/bash
# after a stage-kernel change: stage + render parity tests, then ncu durations of the stage kernels (render chunk and 1024-ray step)
set -u
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build()" > /dev/null 2>&1
timeout 900 python -m pytest tests/test_gpu_stages.py tests/test_gpu_render.py tests/test_gpu_train.py -m gpu -q -p no:cacheprovider -x 2>&1 | tail -3
M=gpu__time_duration.sum
timeout 600 ncu --metrics $M --clock-control none -k regex:"sample_pdf|composite|stratified" -c 24 --csv --log-file gpurun_out/stg_render.csv \
  python bench.py --mode render --steps 1 --warmup 1 --no-gpu-baseline --no-cpu-baseline > /dev/null 2>&1
timeout 600 ncu --metrics $M --clock-control none -k regex:"sample_pdf|composite|stratified|clip_adam" --launch-skip 20 -c 40 --csv --log-file gpurun_out/stg_train.csv \
  python bench.py --mode train --no-graph --steps 1 --warmup 3 --no-gpu-baseline --no-cpu-baseline > /dev/null 2>&1
python profiles/summarize.py launches gpurun_out/stg_render.csv
python profiles/summarize.py launches gpurun_out/stg_train.csv
python bench.py --mode render --no-gpu-baseline --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('render ms %.3f frac %.4f share %.4f' % (d['ms_per_step'], d['roofline']['frac'], d['roofline']['share_of_step']))"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"sample_pdf_kernel" --launch-skip 2 -c 1 -o gpurun_out/prof_pdf_r2 \
  python bench.py --mode render --steps 1 --warmup 1 --no-gpu-baseline --no-cpu-baseline > /dev/null 2>&1
