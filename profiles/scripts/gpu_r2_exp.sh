#!/bin/bash
# Timing experiments on the training kernels (diagnostic builds -DNERF_EXP=bits; results are wrong on purpose):
# bit 0 no arena stores, bit 1 issuers do not wait for weight stages, bit 2 dgrad loads no masks.
set -u
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build()" > /dev/null 2>&1
run() {
  python bench.py --mode train --no-gpu-baseline --no-cpu-baseline --steps 6 --warmup 5 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); t=d['train']
print('$1', 'step %.4f' % t['ms_per_step'], {k:round(v['ms_per_step'],4) for k,v in t['kernels'].items()})"
}
run product
for b in ${EXPS:-1 2 3 4 5}; do NERF_B200_LIB=$PWD/nerf-replication_b200/libnerf_b200_exp$b.so run exp$b; done
