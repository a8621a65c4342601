#!/bin/bash
# round 2, call 30 (2 GPUs): offline loop (symmetric boundary faces refreshed) on 1 and 2 GPUs; OBC 2x1 mismatch dump
mkdir -p gpurun_out/r30
export MOM6ADV_TEST_DUMP=$PWD/gpurun_out/r30
timeout 900 python -m pytest tests/test_offline_loop.py -q -m gpu > gpurun_out/r30/pytest_1gpu.log 2>&1; echo "rc=$?"; tail -4 gpurun_out/r30/pytest_1gpu.log
timeout 1200 python -m pytest tests/test_multi_gpu.py -q -m gpu -k "offline or (open_boundaries and PPM-layout1)" > gpurun_out/r30/pytest_2gpu.log 2>&1; echo "rc=$?"; grep "AssertionError: \[" gpurun_out/r30/pytest_2gpu.log | cut -c1-500; tail -3 gpurun_out/r30/pytest_2gpu.log
ls gpurun_out/r30
