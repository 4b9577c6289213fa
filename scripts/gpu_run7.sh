mkdir -p gpurun_out
timeout 120 python scripts/profile_sweep.py > gpurun_out/ps_plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_sweep.csv python scripts/profile_sweep.py > gpurun_out/ps_ncu.log 2>&1; echo "exit $?"
wc -l gpurun_out/launches_sweep.csv
