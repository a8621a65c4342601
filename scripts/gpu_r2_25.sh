#!/bin/bash
# round 2, call 25 (2 GPUs): multi-GPU tests incl. the hordiff tracers-only pass (both layouts, pull and send/recv)
mkdir -p gpurun_out/r25
timeout 900 python -m pytest tests/test_multi_gpu.py -x -q -m gpu > gpurun_out/r25/pytest_multi_2gpu.log 2>&1; echo "rc=$?"; tail -4 gpurun_out/r25/pytest_multi_2gpu.log
