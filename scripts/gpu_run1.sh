set -x
nvidia-smi --query-gpu=name,memory.total --format=csv
free -g | head -2; nproc
mkdir -p gpurun_out
for f in test_kernels_gpu test_parity_gpu test_properties_gpu; do
  timeout 900 python -m pytest tests/$f.py -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/$f.log 2>&1; echo "exit $f: $?"
  tail -5 gpurun_out/$f.log
done
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit: $?"; tail -3 gpurun_out/smoke.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench1.log 2>&1; echo "bench exit: $?"; tail -2 gpurun_out/bench1.log
