timeout 1200 python -m pytest tests -m gpu -q 2>&1 | tail -3
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r01c_bench_1gpu.json 2> gpurun_out/r01c_bench_1gpu.err; echo "bench rc $?"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r01c_bench_ref.json 2>&1; echo "ref rc $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tc_gemm_kernel -s 42 -c 1 -o gpurun_out/r01c_tcgemm_pipe python scripts/prof_train.py 4096 once > gpurun_out/s30_ncu_nt.log 2>&1; tail -1 gpurun_out/s30_ncu_nt.log
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r01c_train512_launches.csv python scripts/prof_train.py 512 once > gpurun_out/s30_ncu2.log 2>&1
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r01c_launches.csv python bench.py --events 148 --samples 1280 --steps 2 --warmup 3 --no-cpu --train-events-full 0 > gpurun_out/s30_ncu_bench.log 2>&1; tail -1 gpurun_out/s30_ncu_bench.log | cut -c1-200
