export PF_TC_DEBUG=1
timeout 1200 python -m pytest tests -m gpu -q 2>&1 | tail -8
unset PF_TC_DEBUG
for ev in 512 4096; do
timeout 300 python scripts/prof_train.py $ev graph 2>&1 | grep -v "^+"
done
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/s28_train4096_launches.csv python scripts/prof_train.py 4096 once > gpurun_out/s28_ncu.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/s28_train512_launches.csv python scripts/prof_train.py 512 once > gpurun_out/s28_ncu2.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tc_gemm_kernel -s 42 -c 1 -o gpurun_out/r01_tcgemm_pipe python scripts/prof_train.py 4096 once > gpurun_out/s28_ncu_nt.log 2>&1; tail -2 gpurun_out/s28_ncu_nt.log
