#!/bin/bash
# GPU box: frame tests, then the bench frame for several NRC_LIGHT_OCC_SHIFT values (0 = occupancy test off)
timeout 400 python -m pytest tests/test_gpu_frame.py -x -q -m gpu > gpurun_out/occ_test.log 2>&1; echo "rc=$?" >> gpurun_out/occ_test.log
for k in ${SHIFTS:-0 5 4 3}; do
  NRC_LIGHT_OCC_SHIFT=$k timeout 120 python bench.py --steps 20 --warmup 5 --no-fit --no-cpu-baseline --no-fp32-line > gpurun_out/occ_$k.json 2> gpurun_out/occ_$k.err
done
python - <<PY
import json, os
print("".join(open("gpurun_out/occ_test.log").readlines()[-8:]))
for k in os.environ.get("SHIFTS", "0 5 4 3").split():
    try:
        d = json.loads(open("gpurun_out/occ_%s.json" % k).read().strip().splitlines()[-1])
        ks = d.get("kernels", {})
        print(k, round(d["ms_per_step"], 4), {n: round(v.get("ms_per_step", 0), 3) for n, v in ks.items() if "walk" in n or "light" in n})
    except Exception as e:
        print(k, "ERR", e, open("gpurun_out/occ_%s.err" % k).read()[-400:])
PY
