#!/bin/bash
# 2-GPU box: the bench frame on two ranks and the data-parallel bit-identity check
T=${TAG:-r01q}
R="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 200 $R --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 --no-fit --no-cpu-baseline --no-fp32-line > gpurun_out/${T}_bench2.json 2> gpurun_out/${T}_bench2.err; echo "bench rc=$?"
timeout 200 $R --master-port 29512 tests/multi_gpu/check_dp.py > gpurun_out/${T}_check_dp2.log 2>&1; echo "check_dp rc=$?"
tail -3 gpurun_out/${T}_check_dp2.log
python - <<PY
import json
d = json.loads(open("gpurun_out/${T}_bench2.json").read().strip().splitlines()[-1])
print({k: d.get(k) for k in ("value", "ms_per_step", "n_gpus", "e2e", "gpu_launches")}, d["config"]["frame"])
PY
