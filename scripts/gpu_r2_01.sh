set -x
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv
timeout 300 scripts/ubench/issue_model > gpurun_out/r02_issue_model.txt 2>&1; tail -5 gpurun_out/r02_issue_model.txt
timeout 300 scripts/ubench/issue_model ncu > gpurun_out/r02_issue_model_short.txt 2>&1 && \
timeout 600 ncu --metrics smsp__issue_active.avg.per_cycle_active,smsp__inst_executed.sum,sm__inst_executed_pipe_fp64.sum,sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active,smsp__cycles_active.avg,smsp__average_warp_latency_issue_stalled_math_pipe_throttle.pct,smsp__average_warp_latency_issue_stalled_dispatch_stall.pct --clock-control none --csv --log-file gpurun_out/r02_issue_model_ncu.csv scripts/ubench/issue_model ncu > gpurun_out/r02_issue_model_ncu.log 2>&1
tail -3 gpurun_out/r02_issue_model_ncu.log
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 300 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu 2>gpurun_out/r2_01.err | tee gpurun_out/r2_01.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['value'], d['gpu_launches'], {k: round(v,3) for k,v in d['roofline']['class_ms_per_step'].items()})" || tail -5 gpurun_out/r2_01.err
