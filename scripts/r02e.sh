#!/bin/bash
T=${TAG:-r02e}
[ -n "$SKIP_TESTS" ] || timeout 1200 python -m pytest tests -x -q -m gpu > gpurun_out/${T}_tests.log 2>&1; echo "rc=$?" >> gpurun_out/${T}_tests.log
tail -25 gpurun_out/${T}_tests.log
SECONDS=0; timeout 600 python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?"
echo "bench wall ${SECONDS}s"
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/${T}_bench.json").read().strip().splitlines()[-1])
    print(d["ms_per_step"], d["e2e"]["ms_per_step"], d["repeats"]); print({k:round(v["ms_per_step"],4) for k,v in d["kernels"].items()})
    print("parity", d.get("parity"))
    for k in ("config3","config5_n1"):
        print(k, d.get(k))
    f=d["fit_config2"]; print(f["train_step_ms"], f["infer_ms"], f["loss_first_last"], {k:v for k,v in f["kernels"].items()})
    print(d["cpu_baseline"])
except Exception as e:
    print("ERR", e); print(open("gpurun_out/${T}_bench.err").read()[-3000:])
PY
