#!/bin/bash
# GPU call: final tree of the round -- all GPU tests, smoke, the three bench workloads at N=1 and the reference arm
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -3 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_f2f.json 2> gpurun_out/bench_f2f.err; tail -2 gpurun_out/bench_f2f.err; cut -c1-300 gpurun_out/bench_f2f.json
timeout 600 python bench.py --workload tvl1 > gpurun_out/bench_tvl1.json 2> gpurun_out/bench_tvl1.err; cut -c1-300 gpurun_out/bench_tvl1.json
timeout 600 python bench.py --workload batch --no-cpu > gpurun_out/bench_batch_n1.json 2> gpurun_out/bench_batch_n1.err; cut -c1-300 gpurun_out/bench_batch_n1.json
timeout 900 python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; cut -c1-300 gpurun_out/bench_ref.json
