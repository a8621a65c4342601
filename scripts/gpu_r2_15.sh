timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_golden.py -m gpu -x -q 2>&1 | tail -4
for d in 0 2 6; do
timeout 600 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu --no-parity --diags $d 2>gpurun_out/r2_15.err | tee gpurun_out/r2_15_diags$d.json | python -c "
import sys,json
d=json.loads([l for l in sys.stdin if l.startswith('{')][-1])
print('diags', d['tracers_with_ad_x_ad_y'], d['ms_per_step'], d['gpu_launches'], {k: round(v,3) for k,v in d['roofline']['class_ms_per_step'].items() if v})" || tail -3 gpurun_out/r2_15.err
done
