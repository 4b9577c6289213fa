"""Exact-GP fit (loss + backward) with the Gram / Cholesky sweep sharded over the GPUs of one node.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        benchmarks/sharded_gp_bench.py --size 16384 [--size 65536]

Prints one JSON line per N from rank 0: sharded fit time (max over ranks, CUDA events) and, at world size 1,
the fused single-GPU path beside it."""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pinot_b200 import ops, sharded_gp  # noqa: E402


def timed(fn, reps):
    fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        if dist.is_initialized():
            dist.barrier()
        torch.cuda.synchronize()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    t = torch.tensor([sorted(ts)[len(ts) // 2]], device="cuda", dtype=torch.float64)
    if dist.is_initialized():
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", dest="n", type=int, action="append")
    ap.add_argument("--d", type=int, default=128)
    ap.add_argument("--block", type=int, default=512)
    ap.add_argument("--reps", type=int, default=3)
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    for n in (a.n or [16384]):
        gen = torch.Generator().manual_seed(0)
        x = 0.1 * torch.randn(n, a.d, generator=gen)
        y = torch.randn(n, 1, generator=gen)
        r0, r1 = n * rank // world, n * (rank + 1) // world
        xl = x[r0:r1].to(dev).requires_grad_(True)
        yl = y[r0:r1].to(dev)
        sc = torch.zeros(a.d, device=dev, requires_grad=True)
        var = torch.zeros((), device=dev, requires_grad=True)
        ls = torch.tensor(-3.0, device=dev, requires_grad=True)
        box = {}

        def fit():
            for t in (xl, sc, var, ls):
                t.grad = None
            loss = sharded_gp.sharded_exact_gp_loss(xl, yl, sc, var, ls, block=a.block, check=False)
            loss.sum().backward()
            box["loss"] = loss.detach()

        ms = timed(fit, a.reps)
        line = {"what": "exact-GP fit, sharded Gram + Cholesky sweep + gradients", "n": n, "d": a.d, "n_gpus": world,
                "block": a.block, "fit_ms": ms, "loss": float(box["loss"]),
                "info": int(sharded_gp._ShardedExactGPLoss.last_info.item())}
        if world == 1:
            xf = x.to(dev).requires_grad_(True)
            yf = y.to(dev)

            def fused():
                for t in (xf, sc, var, ls):
                    t.grad = None
                box["fused"] = ops.exact_gp_loss(xf, yf, sc, var, ls, check=False)
                box["fused"].sum().backward()

            line["fused_single_gpu_ms"] = timed(fused, a.reps)
            line["fused_loss"] = float(box["fused"])
        if rank == 0:
            print(json.dumps(line), flush=True)
        del xl, yl
        torch.cuda.empty_cache()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
