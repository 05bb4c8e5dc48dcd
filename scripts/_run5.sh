export PF_TC_DEBUG=1
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -40
unset PF_TC_DEBUG
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/s23_train_tc_launches.csv python scripts/prof_train.py 512 once > gpurun_out/s23_ncu.log 2>&1
tail -2 gpurun_out/s23_ncu.log
