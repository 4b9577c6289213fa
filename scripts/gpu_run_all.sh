set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/pytest_all.log 2>&1; echo "pytest exit: $?"; tail -3 gpurun_out/pytest_all.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit: $?"; tail -1 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_full.log 2>gpurun_out/bench_full.err; echo "bench exit: $?"; tail -1 gpurun_out/bench_full.log | cut -c1-300
timeout 600 python benchmarks/kernel_bench.py 2>gpurun_out/kb.err > gpurun_out/kernel_bench_final.jsonl; echo "kb exit $?"; grep node_update gpurun_out/kernel_bench_final.jsonl
