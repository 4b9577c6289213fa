"""Data-parallel plumbing for one process per GPU (torch.distributed; NCCL on GPUs, gloo in CPU tests).

The reference has no distributed code at all.  What shards naturally
on this path is the molecule dimension: molecules are independent graphs, so
each rank embeds and fits its own batch and the only exchange per step is ONE all-reduce of the
flattened gradient buffer (about 65k floats for GraphConv 3x128: latency-bound, hence one bucket).
"""
import os

import torch
import torch.distributed as dist


def init_from_env(backend=None):
    """Initialise the default process group from RANK / WORLD_SIZE / MASTER_* (torchrun).
    Returns (rank, world_size, local_rank).  A single process needs no group."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(local_rank)
            dist.init_process_group(backend=backend, device_id=torch.device("cuda", local_rank))
        else:
            dist.init_process_group(backend=backend)
    return rank, world, local_rank


def world_size():
    return dist.get_world_size() if dist.is_initialized() else 1


def shard_range(n_items, rank, world):
    """Contiguous block partition of ``n_items`` over ``world`` ranks (sizes differ by at most 1)."""
    base, rem = divmod(int(n_items), int(world))
    start = rank * base + min(rank, rem)
    return start, start + base + (1 if rank < rem else 0)


def shard_by_weight(weights, world):
    """Boundaries of a contiguous partition of items with the given weights (e.g. atoms per
    molecule) into ``world`` blocks of near-equal total weight.  Returns world + 1 offsets."""
    w = torch.as_tensor(weights, dtype=torch.float64)
    c = torch.cumsum(w, 0)
    total = float(c[-1]) if w.numel() else 0.0
    cuts = [0]
    for r in range(1, world):
        target = total * r / world
        cuts.append(int(torch.searchsorted(c, torch.tensor(target, dtype=torch.float64)).item()) + 1)
    cuts.append(int(w.numel()))
    for i in range(1, len(cuts)):
        cuts[i] = min(max(cuts[i], cuts[i - 1]), int(w.numel()))
    return cuts


class GradientBucket:
    """One flat fp32 buffer for all gradients of ``params``; ``allreduce()`` sums (or averages) it
    across ranks with a single collective and scatters the result back into ``p.grad``."""

    def __init__(self, params):
        self.params = [p for p in params if p.requires_grad]
        self.numel = sum(p.numel() for p in self.params)
        dev = self.params[0].device if self.params else torch.device("cpu")
        self.flat = torch.zeros(self.numel, dtype=torch.float32, device=dev)

    def allreduce(self, average=False):
        if world_size() == 1:
            return
        grads = []
        for p in self.params:
            if p.grad is None:
                p.grad = torch.zeros_like(p)
            grads.append(p.grad)
        # one gather kernel, one collective, one scatter kernel
        torch.cat([g.reshape(-1) for g in grads], out=self.flat)
        dist.all_reduce(self.flat, op=dist.ReduceOp.SUM)
        if average:
            self.flat.div_(world_size())
        views, off = [], 0
        for p in self.params:
            n = p.numel()
            views.append(self.flat[off:off + n].view_as(p))
            off += n
        torch._foreach_copy_(grads, views)


def broadcast_parameters(module, src=0):
    """Make every rank start from rank ``src``'s weights."""
    if world_size() == 1:
        return
    for t in list(module.parameters()) + list(module.buffers()):
        dist.broadcast(t.data, src=src)


def all_gather_rows(x):
    """Concatenate per-rank row blocks (possibly of different lengths) on every rank: the
    embedding all-gather in front of a replicated / row-sharded exact GP."""
    if world_size() == 1:
        return x
    w = world_size()
    n = torch.tensor([x.shape[0]], device=x.device, dtype=torch.long)
    sizes = [torch.zeros_like(n) for _ in range(w)]
    dist.all_gather(sizes, n)
    sizes = [int(s.item()) for s in sizes]
    m = max(sizes)
    pad = torch.zeros((m,) + tuple(x.shape[1:]), device=x.device, dtype=x.dtype)
    pad[: x.shape[0]] = x
    out = [torch.empty_like(pad) for _ in range(w)]
    dist.all_gather(out, pad)
    return torch.cat([o[:s] for o, s in zip(out, sizes)], dim=0)


def max_over_ranks(value, device):
    t = torch.tensor([float(value)], device=device, dtype=torch.float64)
    if world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
