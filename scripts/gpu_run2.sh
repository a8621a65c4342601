set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -15
python __graft_entry__.py smoke 2>&1 | tail -2
python bench.py --workload OM4_1 --steps 5 --warmup 3 --no-e2e --no-cpu 2>&1 | tail -3
python bench.py --workload OM4_1 --cfl 0.3 --steps 5 --warmup 3 --no-e2e --no-cpu 2>&1 | tail -3
python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu --cfl 0.3 2>&1 | tail -3
