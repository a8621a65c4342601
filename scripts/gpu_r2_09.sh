timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 300 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu 2>gpurun_out/r2_09.err | tee gpurun_out/r2_09_n1.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['value'], d['gpu_launches'], {k: round(v,3) for k,v in d['roofline']['class_ms_per_step'].items() if v})" || tail -5 gpurun_out/r2_09.err
N=2
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 --no-e2e --no-cpu 2>gpurun_out/r2_09_n$N.err | tee gpurun_out/r2_09_n$N.json | python -c "import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(d['n_gpus'], d['ms_per_step'], d['config'].get('layout'))
        for k,v in d['roofline']['class_ms_per_step_by_rank'].items(): print('  %-10s'%k, v)" || tail -5 gpurun_out/r2_09_n$N.err
