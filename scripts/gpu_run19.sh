mkdir -p gpurun_out
timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-fit16k > gpurun_out/bench_e2e.log 2>&1; echo "exit $?"; tail -1 gpurun_out/bench_e2e.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['value'], d['e2e'])"
timeout 600 python -m pytest tests -m gpu -q -x --tb=short -p no:cacheprovider 2>&1 | tail -2
