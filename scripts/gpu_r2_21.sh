#!/bin/bash
# round 2, call 21: hordiff -- padding columns skipped, two-cells-per-thread variant; tests under both, variant timing
mkdir -p gpurun_out/r21
timeout 600 python -m pytest tests/test_hordiff.py -x -q -m gpu > gpurun_out/r21/pytest_v0.log 2>&1; echo "v0 rc=$?"; tail -2 gpurun_out/r21/pytest_v0.log
MOM6ADV_HD_VARIANT=6 timeout 600 python -m pytest tests/test_hordiff.py -x -q -m gpu > gpurun_out/r21/pytest_v6.log 2>&1; echo "v6 rc=$?"; tail -2 gpurun_out/r21/pytest_v6.log
MOM6ADV_HD_VARIANT=2 timeout 600 python -m pytest tests/test_hordiff.py -x -q -m gpu > gpurun_out/r21/pytest_v2.log 2>&1; echo "v2 rc=$?"; tail -2 gpurun_out/r21/pytest_v2.log
timeout 600 python scripts/bench_hordiff.py --variants 0,1,2,3,4,5,6,7,8,9,10,11,12,13,0 > gpurun_out/r21/variants.json 2> gpurun_out/r21/variants.err
echo "variants rc=$?"; cat gpurun_out/r21/variants.json
