export PF_TC_DEBUG=1
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_glue.py -m gpu -q -x 2>&1 | tail -4
unset PF_TC_DEBUG
for ev in 512 4096; do
timeout 300 python scripts/prof_train.py $ev graph 2>&1 | grep -v "^+"
done
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/s29_train4096_launches.csv python scripts/prof_train.py 4096 once > gpurun_out/s29_ncu.log 2>&1
