#!/bin/bash
# GPU call (8 GPUs): the three bench workloads under torchrun on the final tree (FP16 default)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511"
timeout 900 $TR bench.py --gpus 8 > gpurun_out/bench_f2f_n8.json 2> gpurun_out/bench_f2f_n8.err; tail -2 gpurun_out/bench_f2f_n8.err; cut -c1-260 gpurun_out/bench_f2f_n8.json
timeout 600 $TR bench.py --gpus 8 --workload batch --no-cpu > gpurun_out/bench_batch_n8.json 2> gpurun_out/bench_batch_n8.err; tail -2 gpurun_out/bench_batch_n8.err; cut -c1-260 gpurun_out/bench_batch_n8.json
timeout 600 $TR bench.py --gpus 8 --workload tvl1 --no-cpu > gpurun_out/bench_tvl1_n8.json 2> gpurun_out/bench_tvl1_n8.err; cut -c1-260 gpurun_out/bench_tvl1_n8.json
