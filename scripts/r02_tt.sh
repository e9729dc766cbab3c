#!/bin/bash
# train / inference kernel experiments: stand-alone timings, the MLP test files, two frame sizes
for f in ws 512 256; do echo "form $f"; NRC_TRAIN_TC_FORM=$f timeout 120 python scripts/train_once.py 262144 5 2>&1 | tail -1; done
timeout 120 python scripts/train_once.py 10000 5 2>&1 | tail -1
timeout 600 python -m pytest tests/test_gpu_tc.py tests/test_gpu_ops.py tests/test_gpu_frame.py tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -5
[ -n "$SKIP_FRAMES" ] || bash scripts/r02_sweep.sh "A=1" "A=1 -- --config 3" "A=1"
