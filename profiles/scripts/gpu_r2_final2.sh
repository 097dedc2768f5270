#!/bin/bash
# Round 2 evidence run on 1 x B200: full GPU suite, the default bench line, ncu launch lists (render frame, training step)
# with DRAM bytes, ncu --set full of the training step's MLP kernels.
set -u
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build()" > gpurun_out/build_r2z.log 2>&1
# the allocation-free entry point alone in a fresh process first, then the whole suite (which collects it last)
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider > gpurun_out/pytest_zz_alone_r2z.log 2>&1
echo "zz alone exit $?"; tail -3 gpurun_out/pytest_zz_alone_r2z.log
timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/pytest_r2z.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_r2z.log; tail -3 gpurun_out/pytest_r2z.log
timeout 900 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_r2z.log 2>&1; echo "smoke exit $?"; tail -3 gpurun_out/smoke_r2z.log
timeout 600 python bench.py > gpurun_out/bench_r2z.json 2> gpurun_out/bench_r2z.err; echo "bench exit $?"
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_r2z.json 2>/dev/null
M=dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum
timeout 600 ncu --metrics $M --clock-control none -c 400 --csv --log-file gpurun_out/launches_render_r2z.csv \
  python bench.py --mode render --steps 1 --warmup 1 --no-gpu-baseline --no-cpu-baseline > gpurun_out/ncu_render_r2z.log 2>&1
timeout 600 ncu --metrics $M --clock-control none --launch-skip 70 -c 130 --csv --log-file gpurun_out/launches_train_r2z.csv \
  python bench.py --mode train --no-graph --steps 1 --warmup 3 --no-gpu-baseline --no-cpu-baseline > gpurun_out/ncu_train_r2z.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"mlp_tc2|wgrad_heads" --launch-skip 24 -c 8 \
  -o gpurun_out/prof_train_r2z python bench.py --mode train --no-graph --steps 1 --warmup 3 --no-gpu-baseline --no-cpu-baseline \
  > gpurun_out/ncu_full_train_r2z.log 2>&1
ls gpurun_out | head -50
