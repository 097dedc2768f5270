#!/bin/bash
# A/B of the number of weight-producer warps: parity tests on the product, then train + render timings per library
set -u
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build()" > /dev/null 2>&1
timeout 900 python -m pytest tests/test_gpu_mlp.py tests/test_gpu_render.py tests/test_gpu_train.py tests/test_gpu_step.py -m gpu -q -p no:cacheprovider -x -k "not loss_curve" 2>&1 | tail -3
for v in "" ${VARIANTS:-_exp0_prod1 _exp0_prod2}; do
  export NERF_B200_LIB=$PWD/nerf-replication_b200/libnerf_b200$v.so
  python bench.py --mode all --no-gpu-baseline --no-cpu-baseline --steps 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); t=d['train']
print('variant \"$v\": render %.2f ms frac %.4f | train step %.4f' % (d['ms_per_step'], d['roofline']['frac'], t['ms_per_step']), {k:round(v['ms_per_step'],4) for k,v in t['kernels'].items()}, 'c5 %.3f' % d['train_c5']['ms_per_step'])"
done
