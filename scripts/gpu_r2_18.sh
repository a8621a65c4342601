timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 600 python bench.py --workload OM4_1 --steps 10 --warmup 3 --no-e2e 2>gpurun_out/r2_18.err | grep '^{' > gpurun_out/r2_18_om41_n1.json
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --workload OM4_1 --steps 10 --warmup 3 2>gpurun_out/r2_18b.err | grep '^{' > gpurun_out/r2_18_om41_n2.json
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 2>gpurun_out/r2_18c.err | grep '^{' > gpurun_out/r2_18_om4025_n2.json
python - <<'PY'
import json
for f in ("om41_n1","om41_n2","om4025_n2"):
    try:
        d=json.loads(open('gpurun_out/r2_18_%s.json'%f).read())
        print(f, d['n_gpus'], round(d['ms_per_step'],3), '%.4g'%d['value'], d['checksum'].get('value'), d.get('parity_slab',{}).get('result'), d.get('cpu_baseline',{}) and d['cpu_baseline'].get('value'))
    except Exception as e:
        print(f, 'FAILED', e)
PY
