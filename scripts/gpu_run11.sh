W="1440,1080,75,16"
for ch in 0 1 2 4 8; do
  echo "== chunk $ch"
  timeout 300 python bench.py --workload $W --e2e-only --trace --mirror-chunk $ch --e2e-steps 2 2>&1 | grep -v "^$" | cut -c1-400
done
