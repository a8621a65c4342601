timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
nproc
( time timeout 900 python bench.py 2>gpurun_out/r2_11_n1.err > gpurun_out/r2_11_n1.json ) 2>&1 | grep real
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2_11_n1.json') if l.startswith('{')][-1])
print('N1', d['ms_per_step'], d['value'], 'ocean', d.get('ocean_cell_tracer_updates_per_s'), d.get('ocean_fraction'))
print(' checksum', d.get('checksum'))
print(' parity', d.get('parity_slab'))
print(' e2e', {k: d['e2e'][k] for k in ('value','ms_per_step','ms_per_step_stats','steps')} if d.get('e2e') else None)
print(' e2e_res', {k: d['e2e_resident'].get(k) for k in ('value','ms_per_step','ms_per_step_stats','error')} if d.get('e2e_resident') else None)
print(' cpu', d.get('cpu_baseline'))
print(' roofline', d['roofline']['frac'], d['roofline']['path']['frac'], {k: round(v,3) for k,v in d['roofline']['class_ms_per_step'].items() if v})
PY
tail -3 gpurun_out/r2_11_n1.err
N=2
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 2>gpurun_out/r2_11_n$N.err | grep '^{' > gpurun_out/r2_11_n$N.json
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2_11_n2.json').read())
print('N2', d['ms_per_step'], d['value'], d['config']['layout'], d['checksum'], d['halo_exchange'])
PY
tail -3 gpurun_out/r2_11_n2.err
( time timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_11_ref.json ) 2>&1 | grep real
cut -c1-1200 gpurun_out/r2_11_ref.json
