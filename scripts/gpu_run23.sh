timeout 600 python -m pytest tests/test_multi_gpu.py -x -q -k "eight" 2>&1 | tail -3
for n in 8 4; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $n --steps 5 --warmup 3 > gpurun_out/bench_pull_n$n.json 2> gpurun_out/bench_pull_n$n.err
  tail -1 gpurun_out/bench_pull_n$n.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['n_gpus'], d['ms_per_step'], d['value'], d['config']['layout'], {k: round(v,3) for k,v in d['roofline']['class_ms_per_step'].items()})" || tail -5 gpurun_out/bench_pull_n$n.err
done
