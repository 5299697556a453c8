#!/bin/bash
# GPU call: new tests, SM clock under sustained conv launches, TV-L1 single pair vs batch, bench with lookahead flows,
# TV-L1 ncu --set full of a working finest-scale iteration, F2F PSNR parity at 540x960 and 1080p vs the FP32 CPU oracle
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_dncnn_gpu.py -m gpu -x -q -k "run_sequence or cli" 2>&1 | tail -5
timeout 300 python tools/trace_conv64.py 2>&1 | grep -v Warn | tee gpurun_out/trace_conv64.log
for cfg in "1 1" "5 5" "16 16"; do timeout 300 env REPS=2 python tools/perf_tvl1.py $cfg 2>&1 | grep rep | sed "s/^/pairs,slots=$cfg: /"; done | tee gpurun_out/perf_tvl1_slots.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_f2f.json 2> gpurun_out/bench_f2f.err; tail -3 gpurun_out/bench_f2f.err; cut -c1-200 gpurun_out/bench_f2f.json
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'k_iter_strip' --launch-skip 556 -c 2 -o gpurun_out/prof_tvl1 env REPS=0 python tools/perf_tvl1.py 16 16 > gpurun_out/ncu_t.log 2>&1; tail -1 gpurun_out/ncu_t.log
timeout 600 python tools/parity_f2f.py 540 960 3 20 gpurun_out/parity_540.json 2>&1 | grep -v Warn | tail -5
timeout 900 python tools/parity_f2f.py 1080 1920 1 20 gpurun_out/parity_1080.json 2>&1 | grep -v Warn | tail -3
