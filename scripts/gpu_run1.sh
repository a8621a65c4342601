set -x
nproc; free -g | head -2; nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv
python -m pytest tests -m gpu -x -q 2>&1 | tail -5
python __graft_entry__.py smoke 2>&1 | tail -2
python bench.py --steps 3 --warmup 3 > gpurun_out/bench_r01_naive.json 2> gpurun_out/bench_r01_naive.err; tail -c 3000 gpurun_out/bench_r01_naive.json; tail -5 gpurun_out/bench_r01_naive.err
python bench.py --workload OM4_1 --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches_naive.csv python bench.py --workload OM4_1 --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_tracer_sweep -s 6 -c 2 -o gpurun_out/prof_naive python bench.py --workload OM4_1 --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu2.log 2>&1
tail -3 gpurun_out/ncu2.log
