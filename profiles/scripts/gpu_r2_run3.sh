#!/bin/bash
# Round 2, GPU run 3 (1 x B200): PRMT mask layout check (training tests), barrier wait profile of the training kernels.
set -u
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build()" > gpurun_out/build_r2c.log 2>&1
timeout 900 python -m pytest tests/test_gpu_train.py tests/test_gpu_step.py -m gpu -q -p no:cacheprovider -k "not loss_curve" > gpurun_out/pytest_r2c.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_r2c.log; tail -4 gpurun_out/pytest_r2c.log
NERF_B200_LIB=$PWD/nerf-replication_b200/libnerf_b200_waitprof.so python profiles/waitprof.py train > gpurun_out/waitprof_r2c.txt 2>&1
cat gpurun_out/waitprof_r2c.txt
python bench.py --mode train --no-gpu-baseline --no-cpu-baseline --steps 10 --warmup 5 > gpurun_out/train_r2c.json 2>/dev/null
python - <<'PY'
import json
d=json.loads(open("gpurun_out/train_r2c.json").read().strip().splitlines()[-1]); t=d["train"]
print("train dev ms %.4f e2e %.4f" % (t["ms_per_step"], t["e2e"]["ms_per_step"]), {k:round(v["ms_per_step"],4) for k,v in t["kernels"].items()})
PY
