run() {
  MOM6ADV_LIB=$PWD/mom6_b200/variants/libmom6adv_b200_$1.so timeout 300 python bench.py --workload 1440,1080,10,52 --steps 5 --warmup 3 --no-e2e --no-cpu 2>gpurun_out/v_$1.err | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$1', round(d['ms_per_step'],3), {k: round(v,3) for k,v in d['roofline']['class_ms_per_step'].items() if v})" || tail -3 gpurun_out/v_$1.err
}
for v in base y7s y13 y13s y13su4 y13sn12 y13sn5 x5 x6 base; do run $v; done
MOM6ADV_LIB=$PWD/mom6_b200/variants/libmom6adv_b200_y13s.so timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
