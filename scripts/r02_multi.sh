#!/bin/bash
# N-GPU bench as the driver launches it; usage: r02_multi.sh N [extra bench args]
N=$1; shift
T=${TAG:-r02m}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29571 bench.py --gpus $N "$@" > gpurun_out/${T}_bench_n$N.json 2> gpurun_out/${T}_bench_n$N.err; echo "rc=$?"
python - <<PY
import json
try:
    d=json.loads([l for l in open("gpurun_out/${T}_bench_n$N.json").read().strip().splitlines() if l.startswith("{")][-1])
    print("N=$N", d["ms_per_step"], d["value"], "e2e", d["e2e"]["ms_per_step"], d["repeats"]["ms_per_step"])
    print({k:round(v["ms_per_step"],4) for k,v in list(d["kernels"].items())[:9]})
    print("parity", {k:v for k,v in (d.get("parity") or {}).items() if k!="note"})
    for k in ("strong_4k","config5"):
        if k in d:
            x=d[k]; print(k, {a:b for a,b in x.items() if a not in ("parity","workload")}); print(k,"parity", {a:b for a,b in (x.get("parity") or {}).items() if a!="note"})
except Exception as e:
    print("ERR", e); print(open("gpurun_out/${T}_bench_n$N.err").read()[-3000:])
PY
