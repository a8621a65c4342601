timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
W="1440,1080,75,16"
timeout 300 python bench.py --workload $W --e2e-only --trace 2 --mirror-chunk 2 --mirror-ahead 2 --e2e-steps 1 2>&1 | grep -v "^$" | cut -c1-300 | tail -10
for cfg in "1 2" "2 2" "4 2" "2 1" "2 3"; do
  set -- $cfg
  echo "== full chunk $1 ahead $2"
  timeout 600 python bench.py --e2e-only --trace 1 --mirror-chunk $1 --mirror-ahead $2 --e2e-steps 2 2>&1 | grep -v "^$" | cut -c1-330 | tail -2
done
