#!/bin/bash
# GPU call: final FP16-default tree (all GPU tests, smoke, bench) + storage-format A/B (FP16 / BF16 / BF16 with FP16 weights)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -3 gpurun_out/smoke.log
V=frame2frame_b200/lib/variants
echo "== mixed (BF16 activations/gradients, FP16 weights)"
F2F_LIB=$PWD/$V/libf2f_mixed.so timeout 600 python tools/parity_vs_golden.py gpurun_out/c0_mixed.json 2>&1 | grep -v Warn | tail -10
for v in "" $V/libf2f_mixed.so $V/libf2f_bf16.so; do
  echo "== perf ${v:-fp16 default}"
  F2F_LIB=${v:+$PWD/$v} timeout 300 python tools/perf_dncnn.py 1080 1920 5 2>&1 | grep -v Warn
done
timeout 900 python bench.py > gpurun_out/bench_f2f.json 2> gpurun_out/bench_f2f.err; tail -3 gpurun_out/bench_f2f.err; cut -c1-260 gpurun_out/bench_f2f.json
F2F_LIB=$PWD/$V/libf2f_mixed.so timeout 900 python bench.py > gpurun_out/bench_f2f_mixed.json 2> gpurun_out/bench_f2f_mixed.err; cut -c1-260 gpurun_out/bench_f2f_mixed.json
