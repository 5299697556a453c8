#!/bin/bash
# GPU call: CTA-pair variant with one barrier arrival per warp: bit-exactness, role trace, iteration time
mkdir -p gpurun_out
timeout 120 python tools/check_pair.py quick > gpurun_out/check_pair_quick.log 2>&1; echo "quick rc=$?"; tail -10 gpurun_out/check_pair_quick.log | cut -c1-200
F2F_CONV_PAIR=1 timeout 200 python tools/trace_conv64.py > gpurun_out/trace_pair2.log 2>&1; grep -v legacy gpurun_out/trace_pair2.log | tail -12 | cut -c1-330
F2F_CONV_PAIR=1 timeout 200 python tools/perf_dncnn.py 1080 1920 5 2>&1 | grep -v Warn
