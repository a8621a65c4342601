#!/bin/bash
# round 2, call 19: tracer_hordiff row -- GPU parity tests, OM4_025 timing, launch list
mkdir -p gpurun_out/r19
timeout 600 python -m pytest tests/test_hordiff.py -x -q -m gpu > gpurun_out/r19/pytest_hordiff.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r19/pytest_hordiff.log
tail -5 gpurun_out/r19/pytest_hordiff.log
timeout 600 python scripts/bench_hordiff.py --steps 5 --warmup 3 > gpurun_out/r19/hordiff_om4_025.json 2> gpurun_out/r19/hordiff_om4_025.err
echo "bench rc=$?"; cat gpurun_out/r19/hordiff_om4_025.json; tail -3 gpurun_out/r19/hordiff_om4_025.err
timeout 300 python scripts/bench_hordiff.py --workload 1440,1080,75,2 --steps 5 --warmup 3 --no-parity > gpurun_out/r19/hordiff_ntr2.json 2>> gpurun_out/r19/hordiff_om4_025.err
cat gpurun_out/r19/hordiff_ntr2.json
