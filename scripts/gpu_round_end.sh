#!/bin/bash
# GPU box, end of a round: GPU suite, smoke, default bench line, reference arm, ncu launch list and one --set full capture of a frame
T=${TAG:-r02z}
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/${T}_tests.log 2>&1; echo "rc=$?" >> gpurun_out/${T}_tests.log
timeout 100 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; echo "rc=$?" >> gpurun_out/${T}_smoke.log
timeout 600 python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${T}_bench_ref.json 2> gpurun_out/${T}_bench_ref.err
A="--steps 2 --warmup 3 --repeats 1 --no-fit --no-cpu-baseline --no-fp32-line --no-side"
if [ -z "$SKIP_NCU" ]; then
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches.csv python bench.py $A > gpurun_out/${T}_ncu_list.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_train_walk|k_render_walk|k_light|k_infer_tc|k_train_tc|k_composite|k_train_combine|k_render_combine|k_reduce_partials|k_nrc_step|k_mrhe_step" --launch-skip 39 --launch-count 13 -o gpurun_out/prof_${T} python bench.py $A > gpurun_out/${T}_ncu_full.log 2>&1
ncu -i gpurun_out/prof_${T}.ncu-rep --page raw --csv > gpurun_out/${T}_raw.csv 2>/dev/null
python scripts/ncu_summary.py gpurun_out/${T}_raw.csv > gpurun_out/${T}_ncu_summary.md 2>&1
python scripts/ncu_summary.py gpurun_out/${T}_raw.csv --json gpurun_out/${T}_ncu_frame.json "profiles/${T}_ncu_summary.md (ncu --set full, one steady-state launch of each frame kernel)"
fi
python - <<PY
import json
T = "${T}"
for f in ("tests", "smoke"):
    print("".join(open("gpurun_out/%s_%s.log" % (T, f)).readlines()[-4:]))
for f in ("bench", "bench_ref"):
    try:
        d = json.loads(open("gpurun_out/%s_%s.json" % (T, f)).read().strip().splitlines()[-1])
        print(f, {k: d.get(k) for k in ("value", "ms_per_step", "e2e", "gpu_launches", "clocks")})
    except Exception as e:
        print(f, "ERR", e, open("gpurun_out/%s_%s.err" % (T, f)).read()[-600:])
PY
ls -la gpurun_out | tail -12
