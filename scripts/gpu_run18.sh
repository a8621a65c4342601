ARGS="--steps 1 --warmup 3 --no-e2e --no-cpu"
timeout 600 python bench.py $ARGS > gpurun_out/plain18.log 2>&1 && \
timeout 1200 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k "regex:k_x_fast|k_y_ring" -s 6 -c 2 --csv --log-file gpurun_out/traffic_om4_025.csv python bench.py $ARGS > gpurun_out/ncu18.log 2>&1
tail -3 gpurun_out/ncu18.log | cut -c1-300; grep -v "^==" gpurun_out/traffic_om4_025.csv | cut -c1-300 | tail -8
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
