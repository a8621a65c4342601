nvidia-smi -L | wc -l
timeout 1500 python -m pytest tests/test_multi_gpu.py -m gpu -q 2>&1 | tail -4 | tee gpurun_out/r2_17_multi_gpu_tests.log
run() {  # tag workload extra-args
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 10 --warmup 3 --workload $2 $3 2>gpurun_out/r2_17_$1.err | grep '^{' > gpurun_out/r2_17_$1.json
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r2_17_$1.json').read())
    print('$1', d['config']['layout'], round(d['ms_per_step'],3), '%.4g'%d['value'], 'chk', d['checksum'].get('value'), d['halo_exchange'])
    for k,v in d['roofline']['class_ms_per_step_by_rank'].items(): print('   %-10s'%k, v)
except Exception as e:
    print('$1 FAILED', e); print(open('gpurun_out/r2_17_$1.err').read()[-1500:])
PY
}
run om4025_1x8 OM4_025 ""
run om4025_2x4 OM4_025 "--layout 2x4"
run eighth_2x4 eighth "--layout 2x4"
run eighth_1x8 eighth ""
