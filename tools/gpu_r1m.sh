#!/bin/bash
# GPU call: wgrad64 with 8-row tiles / 4 stages / two producer groups -- parity tests, iteration time, launch list
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_dncnn_gpu.py -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 300 python tools/perf_dncnn.py 1080 1920 3 2>&1 | grep -v Warn | tee gpurun_out/perf_dncnn_plain.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_f2f.csv python tools/perf_dncnn.py 1080 1920 1 > gpurun_out/ncu_l.log 2>&1
python tools/launch_summary.py gpurun_out/launches_f2f.csv > gpurun_out/launches_f2f.txt; head -12 gpurun_out/launches_f2f.txt
