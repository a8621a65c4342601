set -x
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
for W in "1440,1080,40,24" OM4_025; do
timeout 600 python bench.py --workload $W --cfl 0.3 --steps 3 --warmup 3 --no-e2e --no-cpu 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['config']['iterations_per_call'], d['roofline']['class_ms_per_step'])"
done
timeout 600 python bench.py --cfl 0.45 --steps 3 --warmup 3 --no-e2e --no-cpu 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['config']['iterations_per_call'], d['roofline']['class_ms_per_step'])"
