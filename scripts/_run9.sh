export PF_TC_DEBUG=1
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_glue.py -m gpu -q -x 2>&1 | tail -25
timeout 300 python scripts/diag_train_tc.py 300 lean_default tc1 2>&1 | grep -v "^  grad t"
unset PF_TC_DEBUG
for ev in 512 4096; do
PF_TRAIN_TC_MIN_ROWS=0 timeout 300 python scripts/prof_train.py $ev graph 2>&1 | grep -v "^+"
PF_TRAIN_TC=0 timeout 300 python scripts/prof_train.py $ev graph 2>&1 | grep -v "^+" | sed 's/^/[fp32 gemms] /'
done
