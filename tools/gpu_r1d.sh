#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
python bench.py --steps 3 --warmup 3 > gpurun_out/bench_f2f.json 2> gpurun_out/bench_f2f.err; tail -3 gpurun_out/bench_f2f.err; cat gpurun_out/bench_f2f.json
