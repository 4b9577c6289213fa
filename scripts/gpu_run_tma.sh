set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py -m gpu -q --tb=short -p no:cacheprovider -k "tma or potrf or nll or exact_gp" > gpurun_out/pytest_tma.log 2>&1; echo "exit: $?"; tail -5 gpurun_out/pytest_tma.log
PINOT_B200_TMA=0 timeout 300 python scripts/debug_sweep.py 2>&1 | grep potrf_ext
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/bench_tma.log 2>&1; echo "bench exit: $?"; grep -o '"exact_gp_fit_n16k": {[^}]*}' gpurun_out/bench_tma.log
PINOT_B200_TMA=0 timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/bench_notma.log 2>&1; echo "bench exit: $?"; grep -o '"exact_gp_fit_n16k": {[^}]*}' gpurun_out/bench_notma.log
