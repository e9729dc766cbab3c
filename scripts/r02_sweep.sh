#!/bin/bash
# usage: r02_sweep.sh "<ENV=.. ENV=..> -- <bench args>" ...   one line per setting (ms/frame, e2e, the slowest kernels)
run() {
  local spec="$1"
  local envs="${spec%%--*}"
  local args=""
  if [[ "$spec" == *" -- "* ]]; then args="${spec#* -- }"; envs="${spec%% -- *}"; fi
  echo "== env[$envs] args[$args]"
  env $envs timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-fp32-line --no-fit --no-side $args 2>gpurun_out/sweep_last.err | tail -1 | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); print('ms/frame %.3f  e2e %.3f' % (d['ms_per_step'], d['e2e']['ms_per_step'])); print({k:round(v['ms_per_step'],3) for k,v in list(d['kernels'].items())[:8]})
except Exception as e:
    print('ERR', e); print(open('gpurun_out/sweep_last.err').read()[-1500:])"
}
for s in "$@"; do run "$s"; done
