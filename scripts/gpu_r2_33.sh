#!/bin/bash
# round 2, call 33: full single-GPU suite after the live-buffer / OBC / offline changes; bench.py default; offline timing
mkdir -p gpurun_out/r33
timeout 1700 python -m pytest tests -m gpu -q > gpurun_out/r33/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r33/pytest_gpu.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r33/bench.json 2> gpurun_out/r33/bench.err; echo "bench rc=$?"; cut -c1-400 gpurun_out/r33/bench.json
timeout 600 python scripts/bench_offline.py --steps 3 --warmup 2 > gpurun_out/r33/offline_om4_025.json 2> gpurun_out/r33/offline.err; echo "offline rc=$?"; cut -c1-700 gpurun_out/r33/offline_om4_025.json
