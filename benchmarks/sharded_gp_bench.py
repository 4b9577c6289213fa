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
    timed.last = ts
    return float(t.item())


def breakdown(x, y, sc, var, ls, rank, world, block, r0, r1):
    """Phases of one fit, each between device synchronisations (so overlap across phases is lost)."""
    import time
    n = x.shape[0]
    out = {}

    def lap(name, fn):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        r = fn()
        torch.cuda.synchronize()
        out[name] = (time.perf_counter() - t0) * 1e3
        return r

    for _ in range(2):
        sw = sharded_gp.ShardedSweep(n, block, rank, world, None, x.device, True, sharded_gp.CudaKernels())
        lap("assemble (Gram blocks)", lambda: sw.assemble("rbf", x, y, sc, var, ls))
        lap("sweep (factor / broadcast / update)", sw.run)
        lap("nll terms", sw.nll_terms)
        alpha = lap("alpha", sw.alpha)
        s = lap("gradient partial products", lambda: sw.grad_partial(False))
        if world > 1:
            def red():
                a0 = 0
                for r in range(world):
                    b0 = n * (r + 1) // world
                    dist.reduce(s[a0:b0], dst=r)
                    a0 = b0
            lap("row-block reduce", red)
        dk = s[r0:r1]

        def gb():
            dk[:, :n].addr_(alpha[r0:r1], alpha, alpha=-0.5)
            return sw.k.gram_bwd_rows("rbf", x[r0:r1].contiguous(), x, sc, var, dk, True, True)
        lap("Gram backward (row block)", gb)
        del sw, s, dk
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", dest="n", type=int, action="append")
    ap.add_argument("--d", type=int, default=128)
    ap.add_argument("--block", type=int, default=512)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--breakdown", action="store_true", help="also time the phases separately (synchronising between them)")
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
                "block": a.block, "fit_ms": ms, "fit_ms_rank0_reps": timed.last, "loss": float(box["loss"]),
                "info": int(sharded_gp._ShardedExactGPLoss.last_info.item())}
        if a.breakdown:
            line["phases_ms"] = breakdown(x.to(dev), y.to(dev).reshape(-1), sc.detach(), var.detach().reshape(1),
                                          ls.detach().reshape(1), rank, world, a.block, r0, r1)
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
