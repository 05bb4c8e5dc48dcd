for ev in 512 4096; do
timeout 300 python scripts/prof_train.py $ev graph 2>&1 | grep -v "^+"
PF_TRAIN_TC=0 timeout 300 python scripts/prof_train.py $ev graph 2>&1 | grep -v "^+" | sed 's/^/[fp32 gemms] /'
done
timeout 300 python scripts/diag_train_tc.py 129 lean_default tc1 2>&1 | grep -v "^  grad t"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/s24_train_tc_launches.csv python scripts/prof_train.py 512 once > gpurun_out/s24_ncu.log 2>&1
tail -2 gpurun_out/s24_ncu.log
