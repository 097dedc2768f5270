#!/bin/bash
# Round 2, GPU run 2 (1 x B200): full parity suite again (loss-curve gates with the fp64 yardstick, whole views vs the CPU
# oracle), then A/B of the training step: CUDA graph on/off x max-shared carve-out of the stage kernels on/off.
set -u
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build()" > gpurun_out/build_r2b.log 2>&1
timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/pytest_r2b.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_r2b.log
tail -8 gpurun_out/pytest_r2b.log
B="python bench.py --mode train --no-gpu-baseline --no-cpu-baseline --steps 10 --warmup 5"
for rep in 1 2; do
  $B > gpurun_out/ab_graph_carve_$rep.json 2>/dev/null
  $B --no-graph > gpurun_out/ab_nograph_carve_$rep.json 2>/dev/null
  NERF_B200_CARVEOUT=0 $B > gpurun_out/ab_graph_nocarve_$rep.json 2>/dev/null
  NERF_B200_CARVEOUT=0 $B --no-graph > gpurun_out/ab_nograph_nocarve_$rep.json 2>/dev/null
done
for f in gpurun_out/ab_*.json; do python - "$f" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); t=d["train"]
print(sys.argv[1].split("/")[-1], "dev ms %.4f e2e ms %.4f other %.4f" % (t["ms_per_step"], t["e2e"]["ms_per_step"], t["other_ms_per_step"]))
PY
done
