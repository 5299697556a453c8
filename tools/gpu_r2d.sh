#!/bin/bash
# GPU call: per-role cycle trace of the convolution kernels, single-CTA vs CTA pair; odd frame shapes
mkdir -p gpurun_out
timeout 200 python tools/trace_conv64.py > gpurun_out/trace_single.log 2>&1; grep -v legacy gpurun_out/trace_single.log | tail -12 | cut -c1-330
F2F_CONV_PAIR=1 timeout 200 python tools/trace_conv64.py > gpurun_out/trace_pair.log 2>&1; grep -v legacy gpurun_out/trace_pair.log | tail -12 | cut -c1-330
timeout 600 python -m pytest tests -m gpu -x -q -k "degenerate" 2>&1 | tail -8
