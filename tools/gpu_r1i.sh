#!/bin/bash
# 2-GPU call: NCCL paths -- default F2F bench (replicas), TV-L1 pair sharding, batch-mode fine-tune (gradient all-reduce)
mkdir -p gpurun_out
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 900 $TR bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/bench_f2f_n$N.json 2> gpurun_out/bench_f2f_n$N.err; tail -2 gpurun_out/bench_f2f_n$N.err; cut -c1-400 gpurun_out/bench_f2f_n$N.json
timeout 900 $TR bench.py --gpus $N --workload batch --steps 2 --warmup 1 > gpurun_out/bench_batch_n$N.json 2> gpurun_out/bench_batch_n$N.err; tail -2 gpurun_out/bench_batch_n$N.err; cut -c1-300 gpurun_out/bench_batch_n$N.json
timeout 900 $TR bench.py --gpus $N --workload tvl1 --steps 3 --warmup 3 > gpurun_out/bench_tvl1_n$N.json 2> gpurun_out/bench_tvl1_n$N.err; tail -2 gpurun_out/bench_tvl1_n$N.err; cut -c1-300 gpurun_out/bench_tvl1_n$N.json
