mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_active_gpu.py -m gpu -q --tb=short -p no:cacheprovider 2>&1 | tail -15
timeout 900 python benchmarks/bo_bench.py --pool 125000 --rounds 2 2>&1 | tail -3
