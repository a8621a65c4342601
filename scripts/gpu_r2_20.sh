#!/bin/bash
# round 2, call 20: hordiff GPU tests again (halo-row comparison fixed), tile/unroll variants, ncu of the iteration kernel
mkdir -p gpurun_out/r20
timeout 600 python -m pytest tests/test_hordiff.py -x -q -m gpu > gpurun_out/r20/pytest_hordiff.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r20/pytest_hordiff.log
tail -4 gpurun_out/r20/pytest_hordiff.log
timeout 600 python scripts/bench_hordiff.py --variants 0,1,2,3,4,5,6,7,8,9,10,11,0 > gpurun_out/r20/variants.json 2> gpurun_out/r20/variants.err
echo "variants rc=$?"; cat gpurun_out/r20/variants.json
CMD="python scripts/bench_hordiff.py --steps 1 --warmup 1 --itts 2 --no-parity"
timeout 300 $CMD > gpurun_out/r20/plain.json 2> gpurun_out/r20/plain.err && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_hd_apply -s 2 -c 1 -o gpurun_out/r20/hd_apply $CMD > gpurun_out/r20/ncu.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/r20/ncu.log
