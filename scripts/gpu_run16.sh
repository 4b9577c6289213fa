mkdir -p gpurun_out
timeout 600 python benchmarks/vgp_bench.py 2>&1 | tail -3
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_full.log 2>&1; echo "bench exit $?"; tail -1 gpurun_out/bench_full.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['value'], d['e2e']['value'], d['exact_gp_fit_n16k'], d['cpu_baseline']['value'], d['gpu_launches'], d['clocks'])"
