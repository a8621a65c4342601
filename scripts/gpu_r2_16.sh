CMD="python bench.py --workload 1440,1080,8,16 --steps 1 --warmup 3 --no-e2e --no-cpu --no-parity --diags 2"
$CMD > gpurun_out/r2_16_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_y_ring -c 4 -o gpurun_out/prof_r02_ydiag $CMD > gpurun_out/r2_16_ncu.log 2>&1
tail -3 gpurun_out/r2_16_ncu.log
ls -la gpurun_out/*.ncu-rep
