mkdir -p gpurun_out
timeout 120 python scripts/debug_tc.py > gpurun_out/debug_tc.log 2>&1; echo "debug_tc exit: $?"; cat gpurun_out/debug_tc.log | tail -40
PINOT_B200_GEMM=ffma timeout 120 python scripts/debug_tc.py > gpurun_out/debug_ffma.log 2>&1; echo "debug_ffma exit: $?"; tail -8 gpurun_out/debug_ffma.log
