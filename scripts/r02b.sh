#!/bin/bash
T=r02b
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/${T}_tests.log 2>&1; echo "rc=$?" >> gpurun_out/${T}_tests.log
tail -30 gpurun_out/${T}_tests.log
timeout 300 python scripts/tc_err.py > gpurun_out/${T}_tc_err.log 2>&1; cat gpurun_out/${T}_tc_err.log | tail -8
timeout 300 python bench.py --no-cpu-baseline --no-fp32-line --no-side > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err
python - <<PY
import json
d=json.loads(open("gpurun_out/${T}_bench.json").read().strip().splitlines()[-1])
print(d["ms_per_step"], d["e2e"]["ms_per_step"]); print({k:round(v["ms_per_step"],4) for k,v in d["kernels"].items()})
f=d["fit_config2"]; print(f["train_step_ms"], f["infer_ms"], f["loss_first_last"], {k:v for k,v in f["kernels"].items()})
PY
