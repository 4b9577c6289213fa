mkdir -p gpurun_out
timeout 900 python benchmarks/kernel_bench.py 2>gpurun_out/kb.err | tee gpurun_out/kernel_bench_v3.jsonl | cut -c1-250
