set -x
python -m pytest tests -m gpu -x -q > gpurun_out/s19_tests.log 2>&1; echo "tests exit $?" >> gpurun_out/s19_tests.log
tail -5 gpurun_out/s19_tests.log
for m in eager raw graph; do timeout 300 python scripts/prof_train.py 512 $m; done > gpurun_out/s19_train.log 2>&1
for m in eager raw graph; do timeout 300 python scripts/prof_train.py 4096 $m; done >> gpurun_out/s19_train.log 2>&1
cat gpurun_out/s19_train.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/s19_train_launches.csv python scripts/prof_train.py 512 once > gpurun_out/s19_ncu.log 2>&1
tail -3 gpurun_out/s19_ncu.log
