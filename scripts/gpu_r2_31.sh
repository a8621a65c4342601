#!/bin/bash
# round 2, call 31 (2 GPUs): offline loop again (closed-edge faces kept); OBC 2x1 under diagnostic variants
mkdir -p gpurun_out/r31
export MOM6ADV_TEST_DUMP=$PWD/gpurun_out/r31
timeout 900 python -m pytest tests/test_offline_loop.py -q -m gpu > gpurun_out/r31/pytest_1gpu.log 2>&1; echo "rc=$?"; tail -2 gpurun_out/r31/pytest_1gpu.log
timeout 600 python -m pytest tests/test_multi_gpu.py -q -m gpu -k "offline" > gpurun_out/r31/pytest_2gpu_offline.log 2>&1; echo "rc=$?"; tail -2 gpurun_out/r31/pytest_2gpu_offline.log
for var in call1_iter1 call1 pull0 yfirst_call1_iter1; do
  MOM6ADV_OBC_VARIANT=$var timeout 300 python -m pytest tests/test_multi_gpu.py -q -m gpu -k "open_boundaries and PPM-layout1" > gpurun_out/r31/obc_$var.log 2>&1; echo "$var rc=$?"
done
for f in gpurun_out/r31/obc_2x1_*; do echo $f $(grep -c "^tracer" $f); done
