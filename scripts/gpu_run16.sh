timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
W="1440,1080,75,16"
timeout 300 python bench.py --workload $W --e2e-only --trace 2 --e2e-steps 1 2>&1 | grep -v "^$" | cut -c1-300 | tail -10
timeout 900 python bench.py --steps 3 --warmup 3 --trace 1 > gpurun_out/bench16.json 2> gpurun_out/bench16.err; tail -c 3000 gpurun_out/bench16.json; tail -3 gpurun_out/bench16.err
