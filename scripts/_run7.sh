export PF_TC_DEBUG=1
timeout 1200 python -m pytest tests -m gpu -q 2>&1 | tail -25
unset PF_TC_DEBUG
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/s25_bench.json 2> gpurun_out/s25_bench.err; echo "bench rc $?"
tail -3 gpurun_out/s25_bench.err
cat gpurun_out/s25_bench.json
