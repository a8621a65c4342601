one() {  # label, env, workload
  env $2 timeout 600 python bench.py --workload $3 --steps 2 --warmup 3 --no-e2e --no-cpu 2>gpurun_out/dbg.err | python -c "
import sys,json
d=json.loads([l for l in sys.stdin if l.startswith('{')][-1])
print('$1', d['ms_per_step'], d['checksum'].get('value'), d['checksum'].get('tracer_last_mod_bc'), d['parity_slab'].get('result'), d['parity_slab'].get('first_mismatch'), d['parity_slab'].get('error'))" || tail -3 gpurun_out/dbg.err
}
two() {
  env $2 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --workload $3 --steps 2 --warmup 3 2>gpurun_out/dbg2.err | python -c "
import sys,json
d=json.loads([l for l in sys.stdin if l.startswith('{')][-1])
print('$1', d['ms_per_step'], d['checksum'].get('value'), d['checksum'].get('tracer_last_mod_bc'))" || tail -3 gpurun_out/dbg2.err
}
one om41 X=1 360,210,75,22
one om4025_8tr X=1 1440,1080,75,8
one om4025_8tr_again X=1 1440,1080,75,8
one om4025_8tr_sparse0 MOM6ADV_OPTS=sparse=0 1440,1080,75,8
one om4025_8tr_fast0 MOM6ADV_OPTS=fast=0 1440,1080,75,8
two n2 X=1 1440,1080,75,8
two n2_dirty0 MOM6ADV_OPTS=dirty=0 1440,1080,75,8
two n2_pull2_0 MOM6ADV_OPTS=pull2=0 1440,1080,75,8
two n2_overlap0 MOM6ADV_OPTS=overlap=0 1440,1080,75,8
two n2_sparse0 MOM6ADV_OPTS=sparse=0 1440,1080,75,8
