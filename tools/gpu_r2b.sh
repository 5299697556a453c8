#!/bin/bash
# GPU call: CLI with host-thread frame I/O; 1080p PSNR parity with the FP16 default; refreshed launch list and
# ncu --set full captures of the dominant kernels on the FP16 build
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -k "cli or run_sequence or golden" 2>&1 | tail -3
timeout 900 python tools/parity_f2f.py 1080 1920 2 20 gpurun_out/parity_1080_fp16.json 2>&1 | grep -v Warn | tail -4
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_f2f_fp16.csv python tools/perf_dncnn.py 1080 1920 1 > gpurun_out/ncu_l.log 2>&1
python tools/launch_summary.py gpurun_out/launches_f2f_fp16.csv > gpurun_out/launches_f2f_fp16.txt; head -24 gpurun_out/launches_f2f_fp16.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_conv64|k_wgrad64' --launch-skip 58 -c 6 -o gpurun_out/prof_conv_fp16 python tools/perf_dncnn.py 1080 1920 1 > gpurun_out/ncu_c.log 2>&1; tail -2 gpurun_out/ncu_c.log
