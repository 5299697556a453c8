#!/bin/bash
# GPU call: parity tests, smoke, bench (both arms), ncu launch lists + full captures (DnCNN convs, TV-L1 iteration)
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -3 gpurun_out/smoke.log
python bench.py --steps 3 --warmup 3 > gpurun_out/bench_f2f.json 2> gpurun_out/bench_f2f.err; tail -3 gpurun_out/bench_f2f.err; cat gpurun_out/bench_f2f.json
python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_f2f_ref.json 2> gpurun_out/bench_f2f_ref.err; cat gpurun_out/bench_f2f_ref.json
python tools/perf_dncnn.py 1080 1920 2 > gpurun_out/perf_dncnn_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/launches_dncnn.csv python tools/perf_dncnn.py 1080 1920 1 > gpurun_out/ncu_d1.log 2>&1
grep -v Warn gpurun_out/perf_dncnn_plain.log
ncu --set full --clock-control none --import-source on -k regex:'k_wgrad64|k_conv64' -s 34 -c 4 -o gpurun_out/prof_dncnn python tools/perf_dncnn.py 1080 1920 1 > gpurun_out/ncu_d2.log 2>&1
export REPS=1
python tools/perf_tvl1.py 4 4 > gpurun_out/perf_tvl1_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 12000 --csv --log-file gpurun_out/launches_tvl1.csv python tools/perf_tvl1.py 4 4 > gpurun_out/ncu_t1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_iter_strip -s 1000 -c 2 -o gpurun_out/prof_iter_strip python tools/perf_tvl1.py 4 4 > gpurun_out/ncu_t2.log 2>&1
cat gpurun_out/perf_tvl1_plain.log
ls -la gpurun_out
