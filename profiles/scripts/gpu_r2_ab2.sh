#!/bin/bash
# alternating A/B of two libraries (A = product, B = $VARIANT): render + train timings, no tests
set -u
for r in 1 2; do for v in "" $VARIANT; do
  export NERF_B200_LIB=$PWD/nerf-replication_b200/libnerf_b200$v.so
  python bench.py --mode all --no-gpu-baseline --no-cpu-baseline --steps 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); t=d['train']
print('variant \"$v\": render %.2f ms frac %.4f clk %s | train step %.4f' % (d['ms_per_step'], d['roofline']['frac'], d['clocks']['sm_mhz'], t['ms_per_step']), {k:round(v['ms_per_step'],4) for k,v in t['kernels'].items()}, 'c5 %.3f' % d['train_c5']['ms_per_step'])"
done; done
