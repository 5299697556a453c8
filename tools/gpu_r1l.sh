#!/bin/bash
# GPU call: DnCNN parity tests after the stats-reduction / SIMT changes, per-kernel launch list, conv trace, bench
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -6 gpurun_out/pytest_gpu.log
timeout 300 python tools/perf_dncnn.py 1080 1920 3 2>&1 | grep -v Warn | tee gpurun_out/perf_dncnn_plain.log
timeout 300 python tools/trace_conv64.py 2>&1 | grep -v Warn | grep -v legacy | tee gpurun_out/trace_conv64.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_f2f.csv python tools/perf_dncnn.py 1080 1920 1 > gpurun_out/ncu_l.log 2>&1
python tools/launch_summary.py gpurun_out/launches_f2f.csv > gpurun_out/launches_f2f.txt; head -16 gpurun_out/launches_f2f.txt
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_f2f.json 2> gpurun_out/bench_f2f.err; tail -3 gpurun_out/bench_f2f.err; cut -c1-200 gpurun_out/bench_f2f.json
