#!/bin/bash
# GPU call: TV-L1 with wave-aware strip height (parity + throughput at 1, 5, 16 pairs), lib API test
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_tvl1_gpu.py tests/test_lib_api.py -m gpu -x -q 2>&1 | tail -3
for cfg in "1 1" "5 5" "16 16"; do timeout 300 env REPS=2 python tools/perf_tvl1.py $cfg 2>&1 | grep rep | sed "s/^/pairs,slots=$cfg: /"; done | tee gpurun_out/perf_tvl1_slots.log
timeout 600 python bench.py --workload tvl1 --steps 3 --warmup 3 > gpurun_out/bench_tvl1.json 2> gpurun_out/bench_tvl1.err; cut -c1-200 gpurun_out/bench_tvl1.json
