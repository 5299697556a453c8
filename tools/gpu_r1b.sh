#!/bin/bash
mkdir -p gpurun_out
python tools/perf_dncnn.py > gpurun_out/perf_dncnn_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/launches_dncnn.csv python tools/perf_dncnn.py 1080 1920 1 > gpurun_out/ncu_d1.log 2>&1
python tools/perf_dncnn.py 1080 1920 1 > gpurun_out/perf_dncnn_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'k_wgrad64|k_conv64' -s 34 -c 4 -o gpurun_out/prof_dncnn python tools/perf_dncnn.py 1080 1920 1 > gpurun_out/ncu_d2.log 2>&1
cat gpurun_out/perf_dncnn_plain.log | grep -v Warn
