#!/bin/bash
# Round 2, GPU run 1 (1 x B200): parity tests, default bench line, ncu launch lists with DRAM bytes for the streaming
# stages (render chunk size and training batch size), one ncu --set full capture of the training step's MLP kernels.
# Usage: gpurun --timeout 2400 -- 'bash profiles/scripts/gpu_r2_run1.sh'
set -u
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build()" > gpurun_out/build_r2a.log 2>&1
timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/pytest_r2a.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_r2a.log
tail -5 gpurun_out/pytest_r2a.log
timeout 600 python bench.py > gpurun_out/bench_r2a.json 2> gpurun_out/bench_r2a.err
echo "bench exit $?"; tail -c 600 gpurun_out/bench_r2a.json
M=dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum
timeout 600 ncu --metrics $M --clock-control none -k regex:"composite|sample_pdf|stratified|get_rays" -c 80 --csv \
  --log-file gpurun_out/stages_render_r2a.csv python bench.py --mode render --steps 1 --warmup 1 --no-gpu-baseline --no-cpu-baseline \
  > gpurun_out/ncu_stages_render.log 2>&1
timeout 600 ncu --metrics $M --clock-control none --launch-skip 60 -c 120 --csv \
  --log-file gpurun_out/stages_train_r2a.csv python bench.py --mode train --no-graph --steps 1 --warmup 3 --no-gpu-baseline --no-cpu-baseline \
  > gpurun_out/ncu_stages_train.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"mlp_tc2|wgrad_heads" --launch-skip 24 -c 8 \
  -o gpurun_out/prof_train_r2a python bench.py --mode train --no-graph --steps 1 --warmup 3 --no-gpu-baseline --no-cpu-baseline \
  > gpurun_out/ncu_full_train.log 2>&1
ls -la gpurun_out | head -40
