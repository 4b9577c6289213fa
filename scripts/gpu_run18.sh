mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l); echo "gpus: $N"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --steps 10 --warmup 3 --no-fit16k > gpurun_out/bench_n$N.log 2>&1; echo "exit bench n$N: $?"; tail -1 gpurun_out/bench_n$N.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['n_gpus'], d['ms_per_step'], d['value'], d['e2e']['value'])"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29522 benchmarks/bo_bench.py --pool 125000 --rounds 2 > gpurun_out/bo_n$N.log 2>&1; echo "exit bo n$N: $?"; tail -1 gpurun_out/bo_n$N.log | cut -c1-600
