#!/bin/bash
# round 2, call 24: full GPU suite with the final hordiff kernel; hordiff timing + launch list + ncu; bench.py default
mkdir -p gpurun_out/r24
timeout 1700 python -m pytest tests -m gpu -x -q > gpurun_out/r24/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r24/pytest_gpu.log
timeout 600 python scripts/bench_hordiff.py --steps 5 --warmup 3 > gpurun_out/r24/hordiff_om4_025.json 2> gpurun_out/r24/hordiff.err; echo "hordiff rc=$?"; cat gpurun_out/r24/hordiff_om4_025.json
timeout 300 python scripts/bench_hordiff.py --workload OM4_1 --steps 5 --warmup 3 > gpurun_out/r24/hordiff_om4_1.json 2>> gpurun_out/r24/hordiff.err; cat gpurun_out/r24/hordiff_om4_1.json
CMD="python scripts/bench_hordiff.py --steps 1 --warmup 1 --itts 2 --no-parity"
timeout 300 $CMD > gpurun_out/r24/plain.json 2> gpurun_out/r24/plain.err && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r24/hordiff_launches.csv $CMD > gpurun_out/r24/ncu_list.log 2>&1
echo "list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_hd_tile -s 2 -c 1 -o gpurun_out/r24/hd_tile $CMD > gpurun_out/r24/ncu_full.log 2>&1
echo "full rc=$?"
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r24/bench.json 2> gpurun_out/r24/bench.err; echo "bench rc=$?"; cut -c1-600 gpurun_out/r24/bench.json
