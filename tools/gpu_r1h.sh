#!/bin/bash
# GPU call: all GPU tests, smoke, bench (F2F default + batch workload)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -3 gpurun_out/smoke.log
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_f2f.json 2> gpurun_out/bench_f2f.err; tail -3 gpurun_out/bench_f2f.err; cat gpurun_out/bench_f2f.json
timeout 600 python bench.py --workload batch --steps 1 --warmup 1 > gpurun_out/bench_batch.json 2> gpurun_out/bench_batch.err; tail -3 gpurun_out/bench_batch.err; cat gpurun_out/bench_batch.json
