W="1440,1080,40,24"
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
timeout 300 python bench.py --workload $W --cfl 0.3 --steps 3 --warmup 3 --no-e2e --no-cpu 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['ms_per_step'], d['roofline']['class_ms_per_step'])"
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/bench10.json 2> gpurun_out/bench10.err; tail -c 3000 gpurun_out/bench10.json; tail -5 gpurun_out/bench10.err
