timeout 900 python -m pytest tests/test_multi_gpu.py -x -q 2>&1 | tail -4
W="1440,270,75,52"
run() { echo "== $1"; env $1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --workload $W --steps 5 --warmup 3 2>gpurun_out/err24.log | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], {k: round(v,3) for k,v in d['roofline']['class_ms_per_step'].items()})" || tail -5 gpurun_out/err24.log; }
run "MOM6ADV_OVERLAP=1"
run "MOM6ADV_OVERLAP=0"
