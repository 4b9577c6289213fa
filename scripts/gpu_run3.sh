set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q --tb=short -p no:cacheprovider > gpurun_out/pytest_all.log 2>&1; echo "exit all: $?"; tail -4 gpurun_out/pytest_all.log
PINOT_B200_LOOKAHEAD=0 timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q --tb=short -p no:cacheprovider -k "potrf or exact_gp or chol" > gpurun_out/pytest_nolook.log 2>&1; echo "exit nolook: $?"; tail -3 gpurun_out/pytest_nolook.log
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_look.log 2>&1; echo "bench exit: $?"; tail -1 gpurun_out/bench_look.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['value'], d['e2e'], d['clocks']); print({k:(round(v['ms_per_call'],3), v['calls_per_step']) for k,v in d['kernel_shares'].items()})"
PINOT_B200_LOOKAHEAD=0 timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/bench_nolook.log 2>&1; tail -1 gpurun_out/bench_nolook.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['value']); print({k:(round(v['ms_per_call'],3), v['calls_per_step']) for k,v in d['kernel_shares'].items()})"
