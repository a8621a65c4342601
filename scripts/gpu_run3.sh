set -x
W="1440,1080,4,52"
python bench.py --workload $W --cfl 0.3 --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/plain3.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_._fast -s 6 -c 2 -o gpurun_out/prof_fast1 python bench.py --workload $W --cfl 0.3 --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu3.log 2>&1
tail -2 gpurun_out/plain3.log
tail -3 gpurun_out/ncu3.log
