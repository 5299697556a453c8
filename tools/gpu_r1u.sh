#!/bin/bash
# N-GPU call (N = 4 or 8): the three workloads under torchrun, short runs -- does the multi-rank path hold up at full width?
mkdir -p gpurun_out
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/bench_f2f_n$N.json 2> gpurun_out/bench_f2f_n$N.err; tail -2 gpurun_out/bench_f2f_n$N.err | cut -c1-300; cut -c1-250 gpurun_out/bench_f2f_n$N.json
timeout 600 $TR bench.py --gpus $N --workload batch --steps 2 --warmup 1 > gpurun_out/bench_batch_n$N.json 2> gpurun_out/bench_batch_n$N.err; tail -2 gpurun_out/bench_batch_n$N.err | cut -c1-300; cut -c1-250 gpurun_out/bench_batch_n$N.json
timeout 600 $TR bench.py --gpus $N --workload tvl1 --steps 2 --warmup 3 > gpurun_out/bench_tvl1_n$N.json 2> gpurun_out/bench_tvl1_n$N.err; tail -2 gpurun_out/bench_tvl1_n$N.err | cut -c1-300; cut -c1-250 gpurun_out/bench_tvl1_n$N.json
