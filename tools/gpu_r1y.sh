#!/bin/bash
# (historical: recorded when BF16 was the default and F2F_STORE_FP16=1 selected FP16; FP16 is now the default and F2F_STORE_BF16=1 selects BF16)
# GPU call: configs[0] (9 fine-tuned frames, 960x540) against the recorded FP32 oracle run: BF16 build vs FP16 build
mkdir -p gpurun_out
python -m frame2frame_b200.build --force > /dev/null
timeout 600 python tools/parity_vs_golden.py gpurun_out/c0_bf16.json 2>&1 | grep -v Warn
F2F_STORE_FP16=1 python -m frame2frame_b200.build --force > /dev/null
timeout 600 python tools/parity_vs_golden.py gpurun_out/c0_fp16.json 2>&1 | grep -v Warn
timeout 600 python -m pytest tests/test_dncnn_gpu.py -m gpu -x -q 2>&1 | tail -2
