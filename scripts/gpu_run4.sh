set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -5
W="1440,1080,40,24"
python bench.py --workload $W --cfl 0.3 --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/plain4.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_x_fast -s 6 -c 1 -o gpurun_out/prof_xfast python bench.py --workload $W --cfl 0.3 --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu4x.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_y_fast -s 6 -c 1 -o gpurun_out/prof_yfast python bench.py --workload $W --cfl 0.3 --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu4y.log 2>&1
tail -1 gpurun_out/plain4.log
python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu --cfl 0.3 2>&1 | tail -1
