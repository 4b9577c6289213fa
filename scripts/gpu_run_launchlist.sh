set -x
mkdir -p gpurun_out
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-fit16k > gpurun_out/bench_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-fit16k > gpurun_out/ncu_run.log 2>&1
echo "ncu exit $?"; tail -1 gpurun_out/bench_plain.log | cut -c1-600; wc -l gpurun_out/launches.csv
