W="1440,1080,75,16"
timeout 300 python bench.py --workload $W --e2e-only --trace 2 --mirror-chunk 2 --mirror-ahead 2 --e2e-steps 1 2>&1 | grep -v "^$" | cut -c1-300 | tail -12
CUDA_DEVICE_MAX_CONNECTIONS=32 timeout 300 python bench.py --workload $W --e2e-only --trace 2 --mirror-chunk 2 --mirror-ahead 2 --e2e-steps 1 2>&1 | grep -v "^$" | cut -c1-300 | tail -12
