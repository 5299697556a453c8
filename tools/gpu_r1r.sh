#!/bin/bash
# GPU call: weight gradient with M = 64 instructions for the third vertical tap -- parity, then A/B timing
mkdir -p gpurun_out
F2F_WGRAD_M64=1 timeout 600 python -m pytest tests/test_dncnn_gpu.py -m gpu -x -q -k "wgrad or backward" 2>&1 | tail -12
for rep in 1 2; do
F2F_WGRAD_M64=0 timeout 300 python tools/perf_dncnn.py 1080 1920 5 2>&1 | grep -v Warn | sed "s/^/M64 off: /"
F2F_WGRAD_M64=1 timeout 300 python tools/perf_dncnn.py 1080 1920 5 2>&1 | grep -v Warn | sed "s/^/M64 on : /"
done | tee gpurun_out/m64_ab.log
