#!/bin/bash
# round 2, call 28 (2 GPUs): offline loop through the plan's group pass and multi-block OBC: single-GPU parity again, then two GPUs
mkdir -p gpurun_out/r28
timeout 900 python -m pytest tests/test_offline_loop.py tests/test_hordiff.py tests/test_gpu_parity.py -x -q -m gpu -k "offline or hordiff or open_boundaries or resident" > gpurun_out/r28/pytest_1gpu.log 2>&1; echo "rc=$?"; tail -4 gpurun_out/r28/pytest_1gpu.log
timeout 1200 python -m pytest tests/test_multi_gpu.py -q -m gpu -k "offline or hordiff or open_boundaries" > gpurun_out/r28/pytest_2gpu.log 2>&1; echo "rc=$?"; tail -15 gpurun_out/r28/pytest_2gpu.log
