#!/bin/bash
# round 2, call 27: offline loop timing at OM4_025 (cfl 2.4: several trips) and OM4_1
mkdir -p gpurun_out/r27
timeout 900 python scripts/bench_offline.py --steps 3 --warmup 2 > gpurun_out/r27/offline_om4_025.json 2> gpurun_out/r27/offline.err; echo "rc=$?"; cat gpurun_out/r27/offline_om4_025.json; tail -3 gpurun_out/r27/offline.err
timeout 300 python scripts/bench_offline.py --workload OM4_1 --steps 3 --warmup 2 > gpurun_out/r27/offline_om4_1.json 2>> gpurun_out/r27/offline.err; cat gpurun_out/r27/offline_om4_1.json
