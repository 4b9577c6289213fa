mkdir -p gpurun_out
timeout 200 python scripts/debug_tc.py 2>&1 | grep -E "wide|k1000|launches"
timeout 900 python -m pytest tests -m gpu -q -x --tb=short -p no:cacheprovider 2>&1 | tail -2
timeout 600 python benchmarks/kernel_bench.py 2>/dev/null | grep -E "sgemm|potrf|exact-GP" | cut -c1-170
PINOT_B200_WIDE=0 timeout 600 python benchmarks/kernel_bench.py 2>/dev/null | grep -E "8192|16384" | cut -c1-170
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_tc.log 2>&1; tail -1 gpurun_out/bench_tc.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['value'], d['e2e']['value'], d['exact_gp_fit_n16k']['value']); print({k:(round(v['ms_per_call'],3), v['calls_per_step']) for k,v in d['kernel_shares'].items()})"
