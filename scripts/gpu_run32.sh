# round-1 re-entry: parity of the reworked streaming kernels (fast=2) and A/B timing against the round-1 kernels (fast=1)
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
for f in 1 2; do
timeout 600 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu --fast $f 2>gpurun_out/ab_fast$f.err | tee gpurun_out/ab_fast$f.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('fast', $f, d['ms_per_step'], d['value'], d['gpu_launches'], {k: round(v,3) for k,v in d['roofline']['class_ms_per_step'].items()})" || tail -5 gpurun_out/ab_fast$f.err
done
