mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x --tb=short -p no:cacheprovider 2>&1 | tail -3
timeout 300 python scripts/debug_tc.py 2>&1 | grep -E "k1000|split-K|k128 " 
timeout 600 python benchmarks/kernel_bench.py 2>/dev/null | grep -E "sgemm|potrf|exact-GP" | cut -c1-200
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-fit16k > gpurun_out/bench_tc.log 2>&1; tail -1 gpurun_out/bench_tc.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['value'], d['e2e']['value']); print({k:(round(v['ms_per_call'],3), v['calls_per_step']) for k,v in d['kernel_shares'].items()})"
