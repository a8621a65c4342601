W="1440,1080,40,24"
run() { echo "== $1"; MOM6ADV_NVCC_EXTRA="$1" python -m mom6_b200.build --force > /dev/null; timeout 300 python bench.py --workload $W --cfl 0.3 --steps 3 --warmup 3 --no-e2e --no-cpu 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['ms_per_step'], d['roofline']['class_ms_per_step'])"; }
run "-DYW=5"
run "-DYW=6"
run "-DYW=7 -DYUNROLL=1"
run "-DYW=7 -DYUNROLL=2"
run ""
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
