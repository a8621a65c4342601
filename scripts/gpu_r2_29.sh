#!/bin/bash
# round 2, call 29 (2 GPUs): diagnostics of the failing 2-GPU offline / OBC cases; single-GPU offline with the sparse hint
mkdir -p gpurun_out/r29
timeout 900 python -m pytest tests/test_offline_loop.py -x -q -m gpu > gpurun_out/r29/pytest_1gpu.log 2>&1; echo "rc=$?"; tail -4 gpurun_out/r29/pytest_1gpu.log
timeout 1200 python -m pytest tests/test_multi_gpu.py -q -m gpu -k "(offline and 1-layout) or (open_boundaries and PPM-layout1)" > gpurun_out/r29/pytest_2gpu.log 2>&1; echo "rc=$?"; grep "AssertionError" gpurun_out/r29/pytest_2gpu.log | cut -c1-700; tail -3 gpurun_out/r29/pytest_2gpu.log
