set -x
export PF_TC_DEBUG=1
timeout 300 python scripts/diag_train_tc.py 300 lean_default 2>&1 | tail -15
timeout 300 python scripts/diag_train_tc.py 1000 lean_stress 2>&1 | tail -15
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_glue.py -m gpu -x -q 2>&1 | tail -15
unset PF_TC_DEBUG
for m in raw graph; do timeout 300 python scripts/prof_train.py 512 $m; done 2>&1 | grep -v "^+"
for m in raw graph; do timeout 300 python scripts/prof_train.py 4096 $m; done 2>&1 | grep -v "^+"
PF_TRAIN_TC=0 timeout 300 python scripts/prof_train.py 512 graph 2>&1 | grep -v "^+"
