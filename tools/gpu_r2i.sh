#!/bin/bash
# GPU call: ncu launch list of the bench command on the final tree (FP16 default)
mkdir -p gpurun_out
timeout 200 python bench.py --steps 1 --warmup 3 --no-cpu --no-tvl1 > gpurun_out/bench_short.json 2> gpurun_out/bench_short.err; echo "plain rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 20000 --csv --log-file gpurun_out/launches_bench_fp16.csv python bench.py --steps 1 --warmup 3 --no-cpu --no-tvl1 > gpurun_out/ncu_b.log 2>&1; echo "ncu rc=$?"
python tools/launch_summary.py gpurun_out/launches_bench_fp16.csv > gpurun_out/launches_bench_fp16.txt; head -14 gpurun_out/launches_bench_fp16.txt
gzip -f gpurun_out/launches_bench_fp16.csv
