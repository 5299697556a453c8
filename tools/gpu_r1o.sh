#!/bin/bash
# 2-GPU call: F2F replicas (lookahead flows), batch-mode fine-tune (NCCL all-reduce), TV-L1 pair sharding; reference arm under torchrun
mkdir -p gpurun_out
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 900 $TR bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_f2f_n$N.json 2> gpurun_out/bench_f2f_n$N.err; tail -2 gpurun_out/bench_f2f_n$N.err; cut -c1-250 gpurun_out/bench_f2f_n$N.json
timeout 900 $TR bench.py --gpus $N --workload batch --steps 2 --warmup 1 > gpurun_out/bench_batch_n$N.json 2> gpurun_out/bench_batch_n$N.err; tail -2 gpurun_out/bench_batch_n$N.err; cut -c1-250 gpurun_out/bench_batch_n$N.json
timeout 900 $TR bench.py --gpus $N --workload tvl1 --steps 3 --warmup 3 > gpurun_out/bench_tvl1_n$N.json 2> gpurun_out/bench_tvl1_n$N.err; tail -2 gpurun_out/bench_tvl1_n$N.err; cut -c1-250 gpurun_out/bench_tvl1_n$N.json
timeout 900 $TR bench.py --gpus $N --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref_n$N.json 2> gpurun_out/bench_ref_n$N.err; tail -2 gpurun_out/bench_ref_n$N.err; cut -c1-400 gpurun_out/bench_ref_n$N.json
timeout 600 python -m pytest tests/test_batch_gpu.py -m gpu -x -q 2>&1 | tail -2
