timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
W="1440,1080,75,16"
for ch in 1 2 4; do
  echo "== chunk $ch"
  timeout 300 python bench.py --workload $W --e2e-only --trace --mirror-chunk $ch --e2e-steps 2 2>&1 | grep -v "^$" | cut -c1-330 | tail -2
done
timeout 900 python bench.py --steps 3 --warmup 3 --trace > gpurun_out/bench12.json 2> gpurun_out/bench12.err; tail -c 3000 gpurun_out/bench12.json; tail -5 gpurun_out/bench12.err
