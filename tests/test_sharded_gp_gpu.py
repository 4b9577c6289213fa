"""Sharded exact-GP objective on the GPU: the blocked algorithm with the real C-ABI kernels against the
single-GPU fused path (world size 1, several column blocks), and on two GPUs over NCCL when the box has them."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from pinot_b200 import ops, sharded_gp
from conftest import rel_err

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _problem(n, d, seed, kind):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(n, d, generator=g)
    y = torch.randn(n, 1, generator=g)
    sc = (0.1 * torch.randn(d if kind == "rbf" else 1, generator=g)) - 1.5
    return x, y, sc, torch.tensor(0.2), torch.tensor(-2.0)


def _single_gpu(n, d, seed, kind, exact):
    ts = [t.to(DEV).requires_grad_(True) for t in _problem(n, d, seed, kind)]
    loss = ops.exact_gp_loss(*ts, kind=kind, exact_logdet=exact)
    (2.0 * loss).sum().backward()
    return loss.detach(), [t.grad for t in ts]


@pytest.mark.parametrize("kind,exact", [("rbf", False), ("matern52", True)])
@pytest.mark.parametrize("n,d,block", [(700, 32, 128), (130, 7, 64), (4500, 128, 512)])
def test_blocked_sweep_matches_fused_single_gpu_path(kind, exact, n, d, block):
    want_loss, want = _single_gpu(n, d, 7, kind, exact)
    ts = [t.to(DEV).requires_grad_(True) for t in _problem(n, d, 7, kind)]
    loss = sharded_gp.sharded_exact_gp_loss(*ts, kind=kind, exact_logdet=exact, block=block)
    (2.0 * loss).sum().backward()
    assert loss.shape == (1, 1)
    assert rel_err(loss, want_loss) < 1e-5
    for got, w in zip(ts[:2], want[:2]):
        assert rel_err(got.grad, w) < 2e-4
    for got, w in zip(ts[2:], want[2:]):         # hyper-parameters: long cancelling sums, 1e-3 (north_star)
        assert rel_err(got.grad, w) < 1e-3


def test_non_positive_definite_matrix_reports_the_global_pivot():
    x, y, sc, var, ls = (t.to(DEV) for t in _problem(300, 8, 1, "rbf"))
    x[200, 3] = float("nan")
    with pytest.raises(RuntimeError, match="not positive-definite"):
        sharded_gp.sharded_exact_gp_loss(x, y, sc, var, ls, block=64)
    loss = sharded_gp.sharded_exact_gp_loss(x, y, sc, var, ls, block=64, check=False)
    assert loss.shape == (1, 1)
    assert int(sharded_gp._ShardedExactGPLoss.last_info.item()) > 0


def test_sharded_head_inside_net_matches_exact_head():
    import copy
    import pinot_b200 as pb
    from pinot_b200 import synthetic
    from pinot_b200.regressors import ExactGaussianProcessRegressor, ShardedExactGaussianProcessRegressor
    torch.manual_seed(0)
    g, y = synthetic.make_batch(300, seed=5)
    g, y = g.to(DEV), y.to(DEV)
    gt, _ = synthetic.make_batch(40, seed=6)
    gt = gt.to(DEV)
    rep = pb.representation.Sequential(pb.representation.dgl_legacy.gn(), [64, "tanh", 64, "tanh"])
    a = pb.Net(rep, output_regressor_class=ExactGaussianProcessRegressor).to(DEV)
    b = pb.Net(copy.deepcopy(rep), output_regressor_class=ShardedExactGaussianProcessRegressor, block=128).to(DEV)
    b.load_state_dict(a.state_dict())
    for net in (a, b):
        with torch.no_grad():
            net.output_regressor.kernel.scale.fill_(-3.0)
    la, lb = a.loss(g, y), b.loss(g, y)
    la.sum().backward()
    lb.sum().backward()
    assert rel_err(lb, la) < 1e-5
    for (name, pa), (_, pb_) in zip(a.named_parameters(), b.named_parameters()):
        assert rel_err(pb_.grad, pa.grad) < 1e-3, name
    da, db = a.condition(gt), b.condition(gt)
    # two routes to the same posterior (a: cached L^-T and GEMMs, b: one sweep over [K; y; K_te,tr]); each
    # is held to 1e-4 of the oracle in test_parity_gpu.py, so they agree to that bar, not to rounding
    assert rel_err(db.mean, da.mean) < 1e-4
    ma, mb = a.condition_marginal(gt), b.condition_marginal(gt)
    assert rel_err(mb.mean, ma.mean) < 1e-5 and rel_err(mb.scale, ma.scale) < 1e-5


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out, n, d, block, kind, exact, cuts):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    x, y, sc, var, ls = _problem(n, d, 7, kind)
    r0, r1 = cuts[rank], cuts[rank + 1]
    dev = torch.device("cuda", rank)
    xl = x[r0:r1].to(dev).requires_grad_(True)
    yl = y[r0:r1].to(dev).requires_grad_(True)
    sc, var, ls = (t.to(dev).requires_grad_(True) for t in (sc, var, ls))
    loss = sharded_gp.sharded_exact_gp_loss(xl, yl, sc, var, ls, kind=kind, exact_logdet=exact, block=block)
    (2.0 * loss).sum().backward()
    hyper = torch.cat([sc.grad.reshape(-1), var.grad.reshape(-1), ls.grad.reshape(-1)])
    dist.all_reduce(hyper)                      # partial sums -> full gradient
    torch.save({"loss": loss.detach().cpu(), "dx": xl.grad.cpu(), "dy": yl.grad.cpu(), "hyper": hyper.cpu()}, out % rank)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
@pytest.mark.parametrize("n,d,block,kind,exact", [(1500, 32, 256, "rbf", False), (4600, 128, 512, "rbf", False),
                                                   (900, 16, 128, "matern52", True)])
def test_two_gpus_over_nccl_match_single_gpu(tmp_path, n, d, block, kind, exact):
    want_loss, want = _single_gpu(n, d, 7, kind, exact)
    cuts = [0, n // 2 + 37, n]
    out = str(tmp_path / "r%d.pt")
    mp.spawn(_worker, args=(2, _free_port(), out, n, d, block, kind, exact, cuts), nprocs=2, join=True)
    res = [torch.load(out % r) for r in range(2)]
    want_hyper = torch.cat([want[2].reshape(-1), want[3].reshape(-1), want[4].reshape(-1)]).cpu()
    for r in range(2):
        assert rel_err(res[r]["loss"], want_loss.cpu()) < 1e-5
        assert rel_err(res[r]["dx"], want[0][cuts[r]:cuts[r + 1]].cpu()) < 2e-4
        assert rel_err(res[r]["dy"], want[1][cuts[r]:cuts[r + 1]].cpu()) < 2e-4
        assert rel_err(res[r]["hyper"], want_hyper) < 1e-3
