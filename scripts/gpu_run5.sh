set -x
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
W="1440,1080,40,24"
timeout 300 python bench.py --workload $W --cfl 0.3 --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/plain5.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_x_fast -s 6 -c 1 -o gpurun_out/prof_xfast2 python bench.py --workload $W --cfl 0.3 --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu5x.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_y_ring -s 6 -c 1 -o gpurun_out/prof_yring python bench.py --workload $W --cfl 0.3 --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu5y.log 2>&1
tail -1 gpurun_out/plain5.log
timeout 600 python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu --cfl 0.3 2>&1 | tail -1
