mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_n2.log 2>&1; echo "exit n2: $?"; tail -3 gpurun_out/bench_n2.log | cut -c1-1200
timeout 300 python bench.py --gpus 1 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n1.log 2>&1; echo "exit n1: $?"; tail -1 gpurun_out/bench_n1.log | cut -c1-400
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 1 --warmup 1 --molecules 600 > gpurun_out/bench_ref_n2.log 2>&1; echo "exit ref n2: $?"; tail -1 gpurun_out/bench_ref_n2.log | cut -c1-300
