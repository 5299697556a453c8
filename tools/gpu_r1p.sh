#!/bin/bash
# GPU call: programmatic dependent launch A/B on one box (iteration time with and without), DnCNN parity tests with it on
mkdir -p gpurun_out
for rep in 1 2; do
F2F_PDL=0 timeout 300 python tools/perf_dncnn.py 1080 1920 5 2>&1 | grep -v Warn | sed 's/^/PDL off: /'
F2F_PDL=1 timeout 300 python tools/perf_dncnn.py 1080 1920 5 2>&1 | grep -v Warn | sed 's/^/PDL on : /'
done | tee gpurun_out/pdl_ab.log
timeout 900 python -m pytest tests/test_dncnn_gpu.py tests/test_batch_gpu.py -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
