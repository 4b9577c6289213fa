mkdir -p gpurun_out
timeout 200 python scripts/debug_sweep.py 2>&1 | tee gpurun_out/debug_sweep_tc.log
PINOT_B200_GEMM=ffma timeout 200 python scripts/debug_sweep.py 2>&1 | tee gpurun_out/debug_sweep_ffma.log
