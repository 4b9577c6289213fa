mkdir -p gpurun_out
for v in 0 1 2 3; do echo "variant $v"; PINOT_B200_MP_VARIANT=$v timeout 300 python benchmarks/kernel_bench.py --only-mp 2>&1 | grep '"H": 128, "norm": "both"\|"H": 32, "norm": "both"' | cut -c1-200; done
timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -q --tb=line -p no:cacheprovider -k "aggregate or readout" 2>&1 | tail -2
