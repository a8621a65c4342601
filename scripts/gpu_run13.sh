timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
W="1440,1080,75,16"
for cfg in "1 2" "2 2" "4 1" "4 2" "2 4"; do
  set -- $cfg
  echo "== chunk $1 ahead $2"
  timeout 300 python bench.py --workload $W --e2e-only --trace --mirror-chunk $1 --mirror-ahead $2 --e2e-steps 2 2>&1 | grep -v "^$" | cut -c1-400 | tail -2
done
timeout 900 python bench.py --e2e-only --trace 2>&1 | cut -c1-400 | tail -4
