#!/bin/bash
# GPU call: train-iteration timing, DnCNN parity tests, ncu full capture of the SIMT first/last-layer kernels
mkdir -p gpurun_out
timeout 300 python tools/perf_dncnn.py 1080 1920 3 2>&1 | grep -v Warn
timeout 300 python -m pytest tests/test_dncnn_gpu.py -m gpu -x -q 2>&1 | tail -2
timeout 300 python tools/perf_dncnn.py 1080 1920 1 > gpurun_out/perf_dncnn_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'k_conv_first|k_conv_last|k_wgrad_first' -c 5 -o gpurun_out/prof_simt python tools/perf_dncnn.py 1080 1920 1 > gpurun_out/ncu_s.log 2>&1
tail -3 gpurun_out/ncu_s.log
