# final evidence of the round: parity, smoke, full bench line, launch list and captures of the final kernels, full-size traffic
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python __graft_entry__.py smoke 2>&1 | tail -1
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_r01_final4.json 2> gpurun_out/bench_r01_final4.err; tail -c 200 gpurun_out/bench_r01_final4.json
W="1440,1080,40,24"
ARGS="--workload $W --steps 2 --warmup 3 --no-e2e --no-cpu"
timeout 300 python bench.py $ARGS > gpurun_out/plain45.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:^k_ --csv --log-file gpurun_out/launches_r01d.csv python bench.py $ARGS > gpurun_out/ncu45a.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_x_fast -s 3 -c 1 -o gpurun_out/prof_r01d_x python bench.py $ARGS > gpurun_out/ncu45x.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_y_ring -s 3 -c 1 -o gpurun_out/prof_r01d_y python bench.py $ARGS > gpurun_out/ncu45y.log 2>&1
FARGS="--steps 1 --warmup 3 --no-e2e --no-cpu"
timeout 1200 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k "regex:k_x_fast|k_y_ring" -s 6 -c 2 --csv --log-file gpurun_out/traffic_om4_025_d.csv python bench.py $FARGS > gpurun_out/ncu45t.log 2>&1
grep -c k_ gpurun_out/launches_r01d.csv
