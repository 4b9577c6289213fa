"""World-size-2 gloo tests (CPU) of the data-parallel host logic."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from pinot_b200 import distributed as pbd


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    r, w, _ = pbd.init_from_env(backend="gloo")
    assert (r, w) == (rank, world)
    torch.manual_seed(100 + rank)
    lin = torch.nn.Sequential(torch.nn.Linear(5, 7), torch.nn.Tanh(), torch.nn.Linear(7, 1))
    pbd.broadcast_parameters(lin, src=0)
    w0 = torch.cat([p.detach().reshape(-1) for p in lin.parameters()])
    x = torch.randn(6, 5)
    lin(x).sum().backward()
    local = torch.cat([p.grad.reshape(-1) for p in lin.parameters()])
    bucket = pbd.GradientBucket(lin.parameters())
    bucket.allreduce(average=False)
    summed = torch.cat([p.grad.reshape(-1) for p in lin.parameters()])
    rows = pbd.all_gather_rows(torch.full((2 + rank, 3), float(rank)))
    mx = pbd.max_over_ranks(10.0 + rank, "cpu")
    torch.save({"w0": w0, "local": local, "summed": summed, "rows": rows, "mx": mx}, out % rank)
    dist.barrier()
    dist.destroy_process_group()


def test_gradient_allreduce_broadcast_and_gather(tmp_path):
    world = 2
    out = str(tmp_path / "r%d.pt")
    mp.spawn(_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    res = [torch.load(out % r) for r in range(world)]
    assert torch.equal(res[0]["w0"], res[1]["w0"]), "broadcast must equalise the weights"
    want = res[0]["local"] + res[1]["local"]
    for r in range(world):
        assert torch.allclose(res[r]["summed"], want, atol=1e-6)
        assert res[r]["rows"].shape == (5, 3)
        assert res[r]["rows"][:2].eq(0).all() and res[r]["rows"][2:].eq(1).all()
        assert res[r]["mx"] == 11.0


def test_sharding_helpers():
    assert [pbd.shard_range(10, r, 3) for r in range(3)] == [(0, 4), (4, 7), (7, 10)]
    assert pbd.shard_range(2, 3, 4) == (2, 2)
    cuts = pbd.shard_by_weight([10, 1, 1, 1, 1, 1, 1, 1, 1, 2], 2)
    assert cuts[0] == 0 and cuts[-1] == 10 and cuts[1] in (1, 2)
    cuts = pbd.shard_by_weight(torch.ones(1000), 8)
    assert [b - a for a, b in zip(cuts[:-1], cuts[1:])] == [125] * 8
