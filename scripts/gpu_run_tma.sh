set -x
mkdir -p gpurun_out
timeout 120 python scripts/debug_tma.py > gpurun_out/debug_tma.log 2>&1; echo "exit: $?"; tail -4 gpurun_out/debug_tma.log
timeout 900 python -m pytest tests/test_kernels_gpu.py -m gpu -q --tb=short -p no:cacheprovider -k "tma or potrf or nll or exact_gp" > gpurun_out/pytest_tma.log 2>&1; echo "exit: $?"; tail -3 gpurun_out/pytest_tma.log
timeout 200 python benchmarks/fit_breakdown.py; timeout 200 python benchmarks/fit_breakdown.py --n 8192;  timeout 200 python benchmarks/fit_breakdown.py --n 4200
