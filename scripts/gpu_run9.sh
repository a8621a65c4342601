W="1440,1080,40,24"
run() { echo "== $1"; MOM6ADV_NVCC_EXTRA="$1" python -m mom6_b200.build --force > /dev/null; timeout 300 python bench.py --workload $W --cfl 0.3 --steps 3 --warmup 3 --no-e2e --no-cpu 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['ms_per_step'], d['roofline']['class_ms_per_step'])"; }
run "-DYUNROLL=1"
run "-DYUNROLL=2"
run ""
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 300 python bench.py --workload $W --cfl 0.3 --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/plain9.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_x_fast -s 3 -c 1 -o gpurun_out/prof_xfast3 python bench.py --workload $W --cfl 0.3 --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu9x.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_y_ring -s 3 -c 1 -o gpurun_out/prof_yring2 python bench.py --workload $W --cfl 0.3 --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu9y.log 2>&1
