#!/bin/bash
# round 2, call 26: offline loop on the device -- GPU parity tests
mkdir -p gpurun_out/r26
timeout 900 python -m pytest tests/test_offline_loop.py -x -q -m gpu > gpurun_out/r26/pytest_offline.log 2>&1; echo "rc=$?"; tail -25 gpurun_out/r26/pytest_offline.log
