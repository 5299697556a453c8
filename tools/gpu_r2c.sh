#!/bin/bash
# GPU call: CTA-pair (cta_group::2) convolution variant against the single-CTA kernels (bit-exactness, timing);
# degenerate frame shapes
mkdir -p gpurun_out
timeout 120 python tools/check_pair.py quick > gpurun_out/check_pair_quick.log 2>&1; echo "quick rc=$?"; tail -40 gpurun_out/check_pair_quick.log | cut -c1-200
if grep -q "ALL OK" gpurun_out/check_pair_quick.log; then
  timeout 180 python tools/check_pair.py > gpurun_out/check_pair.log 2>&1; echo "full rc=$?"; tail -16 gpurun_out/check_pair.log
  echo "== iteration time, pair off / on"
  timeout 200 python tools/perf_dncnn.py 1080 1920 5 2>&1 | grep -v Warn
  F2F_CONV_PAIR=1 timeout 200 python tools/perf_dncnn.py 1080 1920 5 2>&1 | grep -v Warn
fi
timeout 600 python -m pytest tests -m gpu -x -q -k "degenerate" 2>&1 | tail -8
