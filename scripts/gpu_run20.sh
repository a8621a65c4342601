nvidia-smi -L | wc -l
timeout 900 python -m pytest tests/test_multi_gpu.py -x -q 2>&1 | tail -4
for n in 2 4 8; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $n --steps 5 --warmup 3 > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err
  tail -1 gpurun_out/bench_n$n.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['n_gpus'], d['ms_per_step'], d['value'], d['config']['layout'], d['roofline']['class_ms_per_step'])" || tail -5 gpurun_out/bench_n$n.err
done
timeout 600 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['n_gpus'], d['ms_per_step'], d['value'])"
