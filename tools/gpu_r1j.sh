#!/bin/bash
# GPU call: GPU tests, smoke, default bench; ncu launch lists (one train iteration + eval; one TV-L1 batch);
# ncu --set full of the dominant kernels (64->64 conv forward / input gradient / weight gradient, TV-L1 iteration)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -6 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -3 gpurun_out/smoke.log
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_f2f.json 2> gpurun_out/bench_f2f.err; tail -3 gpurun_out/bench_f2f.err; cut -c1-600 gpurun_out/bench_f2f.json
timeout 300 python tools/perf_dncnn.py 1080 1920 3 2>&1 | grep -v Warn | tee gpurun_out/perf_dncnn_plain.log
# launch list: 1 warm iteration + 1 timed iteration + eval forward
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_f2f.csv python tools/perf_dncnn.py 1080 1920 1 > gpurun_out/ncu_l.log 2>&1
python tools/launch_summary.py gpurun_out/launches_f2f.csv > gpurun_out/launches_f2f.txt; head -30 gpurun_out/launches_f2f.txt
# full captures: skip the warm iteration (first ~60 launches), 2 launches of each class
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_conv64|k_wgrad64' --launch-skip 58 -c 6 -o gpurun_out/prof_conv python tools/perf_dncnn.py 1080 1920 1 > gpurun_out/ncu_c.log 2>&1; tail -2 gpurun_out/ncu_c.log
timeout 300 env REPS=1 python tools/perf_tvl1.py 16 16 > gpurun_out/perf_tvl1_plain.log 2>&1; tail -2 gpurun_out/perf_tvl1_plain.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_iter_strip' --launch-skip 700 -c 2 -o gpurun_out/prof_tvl1 env REPS=0 python tools/perf_tvl1.py 16 16 > gpurun_out/ncu_t.log 2>&1; tail -2 gpurun_out/ncu_t.log
ls -la gpurun_out
