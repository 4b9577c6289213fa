mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --tb=line -p no:cacheprovider > gpurun_out/pytest_all.log 2>&1; echo "exit all: $?"; grep -E "^FAILED|^/root|passed|failed|Error" gpurun_out/pytest_all.log | head -40
timeout 200 python scripts/debug_sweep.py 2>&1 | grep -E "potrf|backend" 
PINOT_B200_GEMM=ffma timeout 200 python scripts/debug_sweep.py 2>&1 | grep -E "potrf|backend" 
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_tc.log 2>&1; echo "bench exit: $?"; tail -1 gpurun_out/bench_tc.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['value'], d['e2e']['value'], d['clocks'], d['final_loss']); print({k:(round(v['ms_per_call'],3), v['calls_per_step']) for k,v in d['kernel_shares'].items()})"
