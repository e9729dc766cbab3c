#!/bin/bash
# CTAs of the sharded fused exchange at T = 2^22 (33.5 MB head), N ranks: usage r02_dp_sweep.sh N blocks...
N=$1; shift
for b in "$@"; do
  echo "== NRC_DP_BLOCKS=$b"
  NRC_DP_BLOCKS=$b CHECK_DP_BIG=1 CHECK_DP_ONLY=large timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29614 tests/multi_gpu/check_dp.py 2>&1 | grep "DP-CHECK\|Error"
done
