#!/bin/bash
# GPU call: full GPU test suite, smoke, bench (ours + reference arm), ncu launch list of the bench command itself,
# ncu --set full of the dominant kernels, PSNR parity at 540x960 and 1080p with the final kernels
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -3 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_f2f.json 2> gpurun_out/bench_f2f.err; tail -3 gpurun_out/bench_f2f.err; cut -c1-200 gpurun_out/bench_f2f.json
timeout 600 python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; cut -c1-200 gpurun_out/bench_ref.json
# launch list of the bench command itself (no CPU leg, no secondary TV-L1 block)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 20000 --csv --log-file gpurun_out/launches_bench.csv python bench.py --steps 1 --warmup 3 --no-cpu --no-tvl1 > gpurun_out/ncu_b.log 2>&1
python tools/launch_summary.py gpurun_out/launches_bench.csv > gpurun_out/launches_bench.txt; head -24 gpurun_out/launches_bench.txt; gzip -f gpurun_out/launches_bench.csv
timeout 300 python tools/perf_dncnn.py 1080 1920 1 > gpurun_out/perf_dncnn_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_conv64t|k_wgrad64|k_last|k_first' --launch-skip 62 -c 12 -o gpurun_out/prof_conv python tools/perf_dncnn.py 1080 1920 1 > gpurun_out/ncu_c.log 2>&1; tail -1 gpurun_out/ncu_c.log
timeout 600 python tools/parity_f2f.py 540 960 3 20 gpurun_out/parity_540.json 2>&1 | grep -v Warn | tail -2
timeout 900 python tools/parity_f2f.py 1080 1920 2 20 gpurun_out/parity_1080.json 2>&1 | grep -v Warn | tail -3
