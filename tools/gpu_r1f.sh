#!/bin/bash
# GPU call: DnCNN parity tests, train-iteration timing, ncu launch list of one iteration
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_dncnn_gpu.py tests/test_f2f_step_gpu.py -m gpu -x -q > gpurun_out/pytest_dncnn.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_dncnn.log
tail -4 gpurun_out/pytest_dncnn.log
timeout 300 python tools/perf_dncnn.py 1080 1920 3 > gpurun_out/perf_dncnn_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/launches_dncnn.csv python tools/perf_dncnn.py 1080 1920 1 > gpurun_out/ncu_d1.log 2>&1
grep -v Warn gpurun_out/perf_dncnn_plain.log
python tools/launch_summary.py gpurun_out/launches_dncnn.csv | head -16
