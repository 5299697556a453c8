#!/bin/bash
# (historical: recorded when BF16 was the default and F2F_STORE_FP16=1 selected FP16; FP16 is now the default and F2F_STORE_BF16=1 selects BF16)
# GPU call: FP16 instead of BF16 storage of the DnCNN tensors (same tensor-core rate) -- kernel parity, speed, and the
# PSNR parity of the online loop at 540x960 against the FP32 CPU oracle, next to the BF16 build on the same box
mkdir -p gpurun_out
echo "== BF16 build (default)"; python -m frame2frame_b200.build --force > /dev/null
timeout 600 python -m pytest tests/test_dncnn_gpu.py -m gpu -x -q 2>&1 | tail -2
timeout 300 python tools/perf_dncnn.py 1080 1920 5 2>&1 | grep -v Warn
echo "== FP16 build"; F2F_STORE_FP16=1 python -m frame2frame_b200.build --force > /dev/null
timeout 600 python -m pytest tests/test_dncnn_gpu.py -m gpu -x -q 2>&1 | tail -2
timeout 300 python tools/perf_dncnn.py 1080 1920 5 2>&1 | grep -v Warn
timeout 900 python tools/parity_f2f.py 540 960 3 20 gpurun_out/parity_540_fp16.json 2>&1 | grep -v Warn | tail -4 | cut -c1-200
