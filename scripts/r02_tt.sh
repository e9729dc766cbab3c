python scripts/train_once.py 262144 5 2>&1 | tail -1
NRC_TRAIN_TC_THREADS=256 python scripts/train_once.py 262144 5 2>&1 | tail -1
python scripts/train_once.py 10000 5 2>&1 | tail -1
NRC_TRAIN_TC_THREADS=512 python scripts/train_once.py 10000 5 2>&1 | tail -1
timeout 600 python -m pytest tests/test_gpu_tc.py tests/test_gpu_ops.py tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -5
bash scripts/r02_sweep.sh "A=1" "NRC_TRAIN_TC_THREADS=512" "A=1 -- --config 3" "NRC_TRAIN_TC_THREADS=512 -- --config 3"
