W="1440,1080,40,24"
ARGS="--workload $W --steps 2 --warmup 3 --no-e2e --no-cpu"
timeout 300 python bench.py $ARGS > gpurun_out/plain17.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:^k_ --csv --log-file gpurun_out/launches_r01.csv python bench.py $ARGS > gpurun_out/ncu17a.log 2>&1
tail -2 gpurun_out/plain17.log | cut -c1-1500
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_x_fast -s 3 -c 1 -o gpurun_out/prof_r01_x python bench.py $ARGS > gpurun_out/ncu17x.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_y_ring -s 3 -c 1 -o gpurun_out/prof_r01_y python bench.py $ARGS > gpurun_out/ncu17y.log 2>&1
timeout 600 ncu --set full --clock-control none -k regex:k_mass -s 12 -c 2 -o gpurun_out/prof_r01_mass python bench.py $ARGS > gpurun_out/ncu17m.log 2>&1
ls -la gpurun_out | tail -8
