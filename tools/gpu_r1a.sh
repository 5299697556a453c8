#!/bin/bash
# GPU call: parity tests, conv64 check, TV-L1 bench, ncu launch list + full capture of the iteration kernel
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
python tools/test_conv64.py > gpurun_out/conv64.log 2>&1
python bench.py --steps 3 --warmup 3 > gpurun_out/bench_tvl1.json 2> gpurun_out/bench_tvl1.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_tvl1_ref.json 2> gpurun_out/bench_tvl1_ref.err
export REPS=1
python tools/perf_tvl1.py 4 4 > gpurun_out/perf_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 12000 --csv --log-file gpurun_out/launches_tvl1.csv python tools/perf_tvl1.py 4 4 > gpurun_out/ncu1.log 2>&1
python tools/perf_tvl1.py 4 4 > gpurun_out/perf_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_iter_strip -s 1000 -c 2 -o gpurun_out/prof_iter_strip python tools/perf_tvl1.py 4 4 > gpurun_out/ncu2.log 2>&1
tail -3 gpurun_out/pytest_gpu.log; tail -3 gpurun_out/conv64.log; cat gpurun_out/bench_tvl1.json
