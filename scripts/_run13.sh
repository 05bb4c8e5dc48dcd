timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
for N in 8 2; do
timeout 280 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2952$N bench.py --gpus $N --steps 3 --warmup 3 2> gpurun_out/s31_bench_$N.err | grep "^{" > gpurun_out/r01c_bench_tc_${N}gpu.json
python -c "
import json; d=json.load(open('gpurun_out/r01c_bench_tc_${N}gpu.json')); print($N, d['value'], d['e2e']['value'], d['log_prob']['value'], d['train']['value'], d['train']['ms_per_step'])"
tail -2 gpurun_out/s31_bench_$N.err | cut -c1-300
done
