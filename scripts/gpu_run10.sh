mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_parity_gpu.py -m gpu -q --tb=line -p no:cacheprovider -k "aggregate or net_matches or readout or message" > gpurun_out/pytest_mp.log 2>&1; echo "exit mp tests: $?"; tail -2 gpurun_out/pytest_mp.log
timeout 600 python benchmarks/kernel_bench.py --only-mp 2>&1 | tee gpurun_out/kernel_bench_mp.jsonl | cut -c1-230
timeout 300 python scripts/profile_top_kernels.py > gpurun_out/ptk_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"tgemm_kernel|mp_aggregate_kernel" -o gpurun_out/prof_top python scripts/profile_top_kernels.py > gpurun_out/ptk_ncu.log 2>&1; echo "ncu exit $?"; ls -la gpurun_out/prof_top.ncu-rep
