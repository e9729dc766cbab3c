#!/bin/bash
# GPU box: frame tests, strict-math selftest, then the bench frame at several NRC_RWALK_SPLIT values
timeout 400 python -m pytest tests/test_gpu_frame.py tests/test_gpu_strict_math.py -x -q -m gpu > gpurun_out/split_test.log 2>&1; echo "rc=$?" >> gpurun_out/split_test.log
for k in ${SPLITS:-0 6 8 10}; do
  NRC_RWALK_SPLIT=$k timeout 120 python bench.py --steps 20 --warmup 5 --no-fit --no-cpu-baseline --no-fp32-line > gpurun_out/split_$k.json 2> gpurun_out/split_$k.err
done
python - <<PY
import json, os
print("".join(open("gpurun_out/split_test.log").readlines()[-8:]))
for k in os.environ.get("SPLITS", "0 6 8 10").split():
    try:
        d = json.loads(open("gpurun_out/split_%s.json" % k).read().strip().splitlines()[-1])
        ks = d.get("kernels", {})
        print(k, round(d["ms_per_step"], 4), {n: round(v.get("ms_per_launch", 0), 3) for n, v in ks.items() if "walk" in n or "light" in n})
    except Exception as e:
        print(k, "ERR", e, open("gpurun_out/split_%s.err" % k).read()[-400:])
PY
