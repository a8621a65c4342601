#!/bin/bash
# round 2, call 32 (2 GPUs): OBC 2x1, one trip: are the remaining transports equal too?  (+ multi tests after the live-buffer changes)
mkdir -p gpurun_out/r32
export MOM6ADV_TEST_DUMP=$PWD/gpurun_out/r32
MOM6ADV_OBC_VARIANT=call1_iter1 timeout 300 python -m pytest tests/test_multi_gpu.py -q -m gpu -k "open_boundaries and PPM-layout1" > gpurun_out/r32/obc_call1_iter1.log 2>&1; echo "rc=$?"; grep "AssertionError: \[" gpurun_out/r32/obc_call1_iter1.log | cut -c1-900
unset MOM6ADV_OBC_VARIANT
timeout 900 python -m pytest tests/test_multi_gpu.py -q -m gpu -k "not open_boundaries" > gpurun_out/r32/pytest_2gpu.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/r32/pytest_2gpu.log
timeout 600 python -m pytest tests/test_multi_gpu.py -q -m gpu -k "open_boundaries and layout0" > gpurun_out/r32/pytest_2gpu_obc12.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/r32/pytest_2gpu_obc12.log
ls gpurun_out/r32; head -30 gpurun_out/r32/obc_vhr_call1_iter1_rank0.txt 2>/dev/null
