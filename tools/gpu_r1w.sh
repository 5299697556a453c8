#!/bin/bash
# GPU call: the round's final tree -- all GPU tests, smoke, default bench
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -3 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_f2f.json 2> gpurun_out/bench_f2f.err; tail -3 gpurun_out/bench_f2f.err; cut -c1-220 gpurun_out/bench_f2f.json
timeout 900 python bench.py --workload batch --steps 2 --warmup 1 > gpurun_out/bench_batch_n1.json 2> gpurun_out/bench_batch_n1.err; cut -c1-220 gpurun_out/bench_batch_n1.json
