set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_kernels_gpu.py -m gpu -q -p no:cacheprovider -k "node_update or sgemm" 2>&1 | tail -2
timeout 100 python scripts/debug_node.py | tail -9
timeout 300 python scripts/profile_top_kernels.py > gpurun_out/ptk_plain.log 2>&1 && \
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"tgemm_tma_kernel|tgemm_kernel|mp_aggregate_kernel|split_tf32|node_gemm_kernel" -c 18 -o gpurun_out/prof_top_final python scripts/profile_top_kernels.py > gpurun_out/ptk_ncu.log 2>&1; echo "ncu exit $?"; ls -la gpurun_out/prof_top_final.ncu-rep; tail -3 gpurun_out/ptk_ncu.log
