#!/bin/bash
# weight-image replicas (NERF_B200_WREPLICAS): does spreading the CTA pairs over copies of the stage images help?
set -u
python -c "import __graft_entry__ as g; g.build()" > /dev/null 2>&1
run() {
  python bench.py --mode ${MODE:-train} --no-gpu-baseline --no-cpu-baseline --steps 6 --warmup 4 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
if 'train' in d:
    t=d['train']; print('$1', 'train step %.4f' % t['ms_per_step'], {k:round(v['ms_per_step'],4) for k,v in t['kernels'].items()})
else:
    print('$1', 'render ms %.3f' % d['ms_per_step'], 'frac', round(d['roofline']['frac'],4))"
}
for r in ${REPS:-1 2 4 8 16 37}; do NERF_B200_WREPLICAS=$r run rep$r; done
