#!/bin/bash
# env sweeps of the frame bench: one line per setting (ms/frame, e2e, the six slowest kernels)
run() {
  echo "== $*"
  env "$@" timeout 200 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-fp32-line --no-fit 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('ms/frame %.3f  e2e %.3f' % (d['ms_per_step'], d['e2e']['ms_per_step'])); print({k:round(v['ms_per_step'],3) for k,v in list(d['kernels'].items())[:6]})"
}
for s in "$@"; do run $s; done
