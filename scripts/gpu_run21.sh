# 2 GPUs, per-GPU block = the 8-GPU OM4_025 block (1440 x 135 x 75, 52 tracers)
W="1440,270,75,52"
run() { echo "== $1"; env $1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --workload $W --steps 5 --warmup 3 2>gpurun_out/err21.log | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], {k: round(v,3) for k,v in d['roofline']['class_ms_per_step'].items()})" || tail -5 gpurun_out/err21.log; }
run "X=1"
run "NCCL_MIN_P2P_NCHANNELS=32 NCCL_MAX_P2P_NCHANNELS=32"
run "NCCL_MIN_P2P_NCHANNELS=16 NCCL_MAX_P2P_NCHANNELS=16 NCCL_P2P_NET_CHUNKSIZE=524288"
run "NCCL_DEBUG=INFO"
grep -i "p2p\|channel\|NVLS" gpurun_out/err21.log | head -20
