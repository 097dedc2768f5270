#!/bin/bash
# quick check after a kernel change: training parity tests + the training step's timing (1 x B200, < 1 GPU-minute)
set -u
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build()" > /dev/null 2>&1
timeout 900 python -m pytest tests/test_gpu_train.py tests/test_gpu_step.py -m gpu -q -p no:cacheprovider -k "not loss_curve" 2>&1 | tail -3
for i in 1 2; do
python bench.py --mode train --no-gpu-baseline --no-cpu-baseline --steps 10 --warmup 5 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); t=d['train']
print('train step %.4f e2e %.4f' % (t['ms_per_step'], t['e2e']['ms_per_step']), {k:round(v['ms_per_step'],4) for k,v in t['kernels'].items()}, 'other %.4f' % t['other_ms_per_step'], 'non-mlp (plain launches) %.4f ms = %.3f of the step' % (t['non_mlp_ms_per_step_plain_launches'], t['non_mlp_share_plain_launches']), 'hbm frac %.3f' % t['hbm']['frac'])"
done
