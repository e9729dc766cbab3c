#!/bin/bash
run() {
  echo "== $*"
  env "$@" timeout 200 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-fp32-line --no-fit 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('ms/frame %.3f  e2e %.3f' % (d['ms_per_step'], d['e2e']['ms_per_step'])); print({k:round(v['ms_per_step'],3) for k,v in list(d['kernels'].items())[:6]})"
}
run NRC_TRAIN_SMS=0
run NRC_TRAIN_SMS=12
run NRC_TRAIN_SMS=16
run NRC_TRAIN_SMS=24
run NRC_TRAIN_SMS=32
run NRC_TRAIN_SMS=40
