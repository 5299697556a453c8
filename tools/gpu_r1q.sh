#!/bin/bash
# GPU call: which part of programmatic dependent launch helps / hurts (bit 0 attribute on big kernels, 1 on small, 2 early trigger in big, 3 in small)
mkdir -p gpurun_out
for m in 0 1 3 5 7 11 15 0 2 10 9; do
F2F_PDL=$m timeout 300 python tools/perf_dncnn.py 1080 1920 5 2>&1 | grep -v Warn | sed "s/^/PDL mask $m: /"
done | tee gpurun_out/pdl_ab.log
