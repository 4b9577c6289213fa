set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q --tb=short -p no:cacheprovider > gpurun_out/pytest_all.log 2>&1; echo "exit all: $?"; tail -4 gpurun_out/pytest_all.log
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/pytest_zz_alone.log 2>&1; echo "exit zz alone: $?"; tail -4 gpurun_out/pytest_zz_alone.log
timeout 900 python benchmarks/kernel_bench.py > gpurun_out/kernel_bench.jsonl 2> gpurun_out/kernel_bench.err; echo "kbench exit: $?"; cat gpurun_out/kernel_bench.jsonl
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/bench_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/launches.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu_run.log 2>&1; echo "ncu exit: $?"
tail -c 600 gpurun_out/bench_plain.log
wc -l gpurun_out/launches.csv
