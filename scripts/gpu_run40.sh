# multi-GPU: bitwise layout invariance with the reworked kernels (2 GPUs) and the strong-scaling point
timeout 900 python -m pytest tests/test_multi_gpu.py -x -q -k "two" 2>&1 | tail -3
n=2
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $n --steps 5 --warmup 3 > gpurun_out/scale_n$n.json 2> gpurun_out/scale_n$n.err
tail -1 gpurun_out/scale_n$n.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['n_gpus'], d['ms_per_step'], d['value'], {k: round(v,3) for k,v in d['roofline']['class_ms_per_step'].items()})" || tail -5 gpurun_out/scale_n$n.err
