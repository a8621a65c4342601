# evidence for the reworked streaming kernels: parity, full bench line, launch list, full captures, full-size DRAM traffic
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_r01_final2.json 2> gpurun_out/bench_r01_final2.err; tail -c 300 gpurun_out/bench_r01_final2.json
W="1440,1080,40,24"
ARGS="--workload $W --steps 2 --warmup 3 --no-e2e --no-cpu"
timeout 300 python bench.py $ARGS > gpurun_out/plain38.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:^k_ --csv --log-file gpurun_out/launches_r01c.csv python bench.py $ARGS > gpurun_out/ncu38a.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_x_fast -s 3 -c 1 -o gpurun_out/prof_r01c_x python bench.py $ARGS > gpurun_out/ncu38x.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_y_ring -s 3 -c 1 -o gpurun_out/prof_r01c_y python bench.py $ARGS > gpurun_out/ncu38y.log 2>&1
FARGS="--steps 1 --warmup 3 --no-e2e --no-cpu"
timeout 1200 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k "regex:k_x_fast|k_y_ring" -s 6 -c 2 --csv --log-file gpurun_out/traffic_om4_025_c.csv python bench.py $FARGS > gpurun_out/ncu38t.log 2>&1
ls -la gpurun_out | tail -8
