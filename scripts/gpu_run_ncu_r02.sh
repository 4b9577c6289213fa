# Round-2 evidence: bench launch list, then `ncu --set full` of the Gram and sweep kernels.  Every ncu command runs
# only after the same command has exited 0 without ncu.
set -x
mkdir -p gpurun_out
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-fit16k --no-extras > gpurun_out/r02_bench_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r02_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-fit16k --no-c1 --no-extras > gpurun_out/r02_ncu_run.log 2>&1
echo "launch list exit $?"; wc -l gpurun_out/r02_launches.csv
timeout 300 python scripts/profile_r02_kernels.py gram > gpurun_out/r02_prof_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"gram_tma_kernel" -c 14 -f -o gpurun_out/r02_gram python scripts/profile_r02_kernels.py gram > gpurun_out/r02_gram_ncu.log 2>&1
echo "gram ncu exit $?"
timeout 300 python scripts/profile_r02_kernels.py sweep >> gpurun_out/r02_prof_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"tgemm_kernel|chol_diag_kernel" --launch-skip 67 -c 8 -f -o gpurun_out/r02_sweep python scripts/profile_r02_kernels.py sweep > gpurun_out/r02_sweep_ncu.log 2>&1
echo "sweep ncu exit $?"
ls -la gpurun_out/*.ncu-rep
