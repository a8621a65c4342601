timeout 300 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 300 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu 2>gpurun_out/b46.err | tee gpurun_out/b46.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['value'], d['gpu_launches'], {k: round(v,3) for k,v in d['roofline']['class_ms_per_step'].items()})" || tail -5 gpurun_out/b46.err
