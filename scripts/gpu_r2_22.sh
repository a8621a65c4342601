#!/bin/bash
# round 2, call 22: what bounds k_hd_apply -- ncu of three variants (pairs U=2; U=1 at full occupancy; 64x2 tiles)
mkdir -p gpurun_out/r22
CMD="python scripts/bench_hordiff.py --steps 1 --warmup 1 --itts 2 --no-parity"
for v in 6 13 5; do
  export MOM6ADV_HD_VARIANT=$v
  timeout 300 $CMD > gpurun_out/r22/plain_v$v.json 2> gpurun_out/r22/plain_v$v.err && \
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_hd_apply -s 2 -c 1 -o gpurun_out/r22/hd_v$v $CMD > gpurun_out/r22/ncu_v$v.log 2>&1
  echo "v$v ncu rc=$?"
done
