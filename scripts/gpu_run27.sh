# round-1 final evidence: full bench line, launch list, full captures of the two sweeps, full-size DRAM traffic
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_r01_final.json 2> gpurun_out/bench_r01_final.err; tail -c 400 gpurun_out/bench_r01_final.json
W="1440,1080,40,24"
ARGS="--workload $W --steps 2 --warmup 3 --no-e2e --no-cpu"
timeout 300 python bench.py $ARGS > gpurun_out/plain27.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:^k_ --csv --log-file gpurun_out/launches_r01.csv python bench.py $ARGS > gpurun_out/ncu27a.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_x_fast -s 3 -c 1 -o gpurun_out/prof_r01_x python bench.py $ARGS > gpurun_out/ncu27x.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_y_ring -s 3 -c 1 -o gpurun_out/prof_r01_y python bench.py $ARGS > gpurun_out/ncu27y.log 2>&1
timeout 600 ncu --set full --clock-control none -k regex:k_mass -s 0 -c 4 -o gpurun_out/prof_r01_mass python bench.py $ARGS > gpurun_out/ncu27m.log 2>&1
FARGS="--steps 1 --warmup 3 --no-e2e --no-cpu"
timeout 1200 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k "regex:k_x_fast|k_y_ring" -s 6 -c 2 --csv --log-file gpurun_out/traffic_om4_025.csv python bench.py $FARGS > gpurun_out/ncu27t.log 2>&1
ls -la gpurun_out | tail -6
