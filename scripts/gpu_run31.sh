W="1440,1080,40,24"
ARGS="--workload $W --steps 2 --warmup 3 --no-e2e --no-cpu"
timeout 300 python bench.py $ARGS > gpurun_out/plain31.log 2>&1 && \
timeout 600 ncu --set full --clock-control none -k "regex:k_mass_flux_scan|k_mass_update_list|k_flags_begin|k_sparse" -s 8 -c 8 -o gpurun_out/prof_scan python bench.py $ARGS > gpurun_out/ncu31.log 2>&1
ls -la gpurun_out/prof_scan.ncu-rep
