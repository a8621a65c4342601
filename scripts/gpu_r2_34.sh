#!/bin/bash
# round 2, call 34 (2 GPUs): open boundaries on two GPUs against the per-PE reference emulation (both layouts, 4 cases each)
mkdir -p gpurun_out/r34
timeout 1200 python -m pytest tests/test_multi_gpu.py -q -m gpu -k "open_boundaries" > gpurun_out/r34/pytest_2gpu_obc.log 2>&1; echo "rc=$?"; grep "AssertionError: \[" gpurun_out/r34/pytest_2gpu_obc.log | cut -c1-600; tail -3 gpurun_out/r34/pytest_2gpu_obc.log
