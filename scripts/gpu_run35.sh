# launch list + full captures of the two sweeps with the reworked kernels (reduced level/tracer count so that ncu replays fit)
W="1440,1080,40,24"
ARGS="--workload $W --steps 2 --warmup 3 --no-e2e --no-cpu"
timeout 300 python bench.py $ARGS > gpurun_out/plain35.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:^k_ --csv --log-file gpurun_out/launches_r01b.csv python bench.py $ARGS > gpurun_out/ncu35a.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_x_fast -s 3 -c 1 -o gpurun_out/prof_r01b_x python bench.py $ARGS > gpurun_out/ncu35x.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_y_ring -s 3 -c 1 -o gpurun_out/prof_r01b_y python bench.py $ARGS > gpurun_out/ncu35y.log 2>&1
tail -2 gpurun_out/plain35.log | cut -c1-600
ls -la gpurun_out | tail -8
