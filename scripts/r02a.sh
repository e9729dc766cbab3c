#!/bin/bash
# first GPU call of round 2: GPU suite on the multi-set train trace, timeline, walks-in-flight sweep at 1080p and 720p
T=r02a
timeout 600 python -m pytest tests -x -q -m gpu > gpurun_out/${T}_tests.log 2>&1; echo "rc=$?" >> gpurun_out/${T}_tests.log
tail -5 gpurun_out/${T}_tests.log
NRC_BENCH_TIMELINE=1 timeout 300 python bench.py --no-cpu-baseline --no-fit --no-fp32-line --no-side > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err
tail -c 600 gpurun_out/${T}_bench.json; grep TIMELINE gpurun_out/${T}_bench.err | head -60
scripts/r02_sweep.sh "NRC_TRAIN_WALK=1" "NRC_TRAIN_WALK=2" "NRC_TRAIN_WALK=3" "NRC_TRAIN_WALK=1 -- --config 3" "NRC_TRAIN_WALK=2 -- --config 3" "NRC_TRAIN_WALK=3 -- --config 3" "NRC_TRAIN_WALK=3 NRC_E2E_DEPTH=3 -- --config 3" 2>&1 | tee gpurun_out/${T}_sweep.log
