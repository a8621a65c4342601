run() { echo "== $1"; env $1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 8 --steps 5 --warmup 3 2>gpurun_out/err25.log | tail -1 | tee gpurun_out/bench25_$2.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['value'], {k: round(v,3) for k,v in d['roofline']['class_ms_per_step'].items()})" || tail -5 gpurun_out/err25.log; }
run "MOM6ADV_OVERLAP=1" ov1
run "MOM6ADV_OVERLAP=0" ov0
