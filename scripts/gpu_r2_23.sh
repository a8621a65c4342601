#!/bin/bash
# round 2, call 23: hordiff shared-memory-tiled variants (cp.async pipeline): tests under one, timing of all
mkdir -p gpurun_out/r23
MOM6ADV_HD_VARIANT=20 timeout 600 python -m pytest tests/test_hordiff.py -x -q -m gpu > gpurun_out/r23/pytest_v20.log 2>&1; echo "v20 rc=$?"; tail -2 gpurun_out/r23/pytest_v20.log
MOM6ADV_HD_VARIANT=23 timeout 600 python -m pytest tests/test_hordiff.py -x -q -m gpu > gpurun_out/r23/pytest_v23.log 2>&1; echo "v23 rc=$?"; tail -2 gpurun_out/r23/pytest_v23.log
timeout 600 python scripts/bench_hordiff.py --variants 6,20,21,22,23,24,25,26,27,28,6 > gpurun_out/r23/variants.json 2> gpurun_out/r23/variants.err
echo "variants rc=$?"; cat gpurun_out/r23/variants.json
