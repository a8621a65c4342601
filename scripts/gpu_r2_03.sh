timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
run() {
  MOM6ADV_LIB=$2 timeout 300 python bench.py --workload 1440,1080,10,52 --steps 5 --warmup 3 --no-e2e --no-cpu 2>gpurun_out/v_$1.err | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$1', round(d['ms_per_step'],3), {k: round(v,3) for k,v in d['roofline']['class_ms_per_step'].items() if v})" || tail -3 gpurun_out/v_$1.err
}
run march $PWD/mom6_b200/libmom6adv_b200.so
for v in "$@"; do run $v $PWD/mom6_b200/variants/libmom6adv_b200_$v.so; done
