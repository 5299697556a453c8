#!/bin/bash
# GPU call: FP16 storage as the default + gradient scaling -- all GPU tests, smoke, configs[0] parity vs the recorded oracle,
# iteration time, default bench
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -3 gpurun_out/smoke.log
timeout 600 python tools/parity_vs_golden.py gpurun_out/c0_fp16.json 2>&1 | grep -v Warn | tail -3
timeout 300 python tools/perf_dncnn.py 1080 1920 5 2>&1 | grep -v Warn
timeout 900 python bench.py > gpurun_out/bench_f2f.json 2> gpurun_out/bench_f2f.err; tail -3 gpurun_out/bench_f2f.err; cut -c1-260 gpurun_out/bench_f2f.json
