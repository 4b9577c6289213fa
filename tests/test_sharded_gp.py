"""World-size-2 gloo test (CPU) of the sharded exact-GP orchestration: column-block ownership, panel
broadcasts, the row-block reduction of the gradient and the partial-sum convention for the
hyper-parameters.  The per-rank kernels are replaced by a torch/oracle stand-in with the same interface
(the C-ABI kernels themselves are covered on the GPU in test_sharded_gp_gpu.py)."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import gp_ref
from pinot_b200 import sharded_gp


class OracleKernels:
    """CPU stand-in for sharded_gp.CudaKernels (test infrastructure only)."""

    @staticmethod
    def _k(kind, a, b, sc, var):
        f = gp_ref.rbf if kind == "rbf" else gp_ref.matern52_safe
        return f(a, b, sc if kind == "rbf" else sc.reshape(()), var.reshape(()))

    def gram_block(self, kind, x_rows, x_cols, sc, var, out, ld, ls):
        k = self._k(kind, x_rows, x_cols, sc, var)
        w = x_cols.shape[0]
        k[:w] += (sharded_gp.EPSILON + torch.exp(ls)) * torch.eye(w)
        out[: k.shape[0], :w] = k

    def factor_panel(self, blk, w, n_dense, identity):
        a = torch.tril(blk[:w, :w]).double()
        a = a + torch.tril(a, -1).t()
        l, info = torch.linalg.cholesky_ex(a)
        rows = w + n_dense + (w if identity else 0)
        x = blk[w:rows, :w].double()
        blk[w:rows, :w] = torch.linalg.solve_triangular(l, x.t(), upper=False).t().float()
        blk[:w, :w] = torch.tril(l).float() + torch.triu(blk[:w, :w], 1)
        return torch.diagonal(l).float().contiguous(), info.to(torch.int32).reshape(1)

    def use_presplit(self, n):
        return False

    def gemm_nt(self, m, n, k, alpha, a, b, beta, c):
        c[:m, :n] = beta * c[:m, :n] + alpha * (a[:m, :k] @ b[:n, :k].t())

    def nll_terms(self, diag, z, w, out3):
        zz = (z[:w] * z[:w]).sum()
        c = sharded_gp.half_log_2pi_n(w)
        out3[0] = 0.5 * zz + diag.sum() + c
        out3[1] = 0.5 * zz + torch.log(diag).sum() + c
        out3[2] = zz

    def gemv(self, m, rows, cols, v, y):
        y[:rows] = m[:rows, :cols] @ v[:cols]

    def gram_bwd_rows(self, kind, x_rows, x_all, sc, var, dk_rows, need_scale, need_var):
        # "row-side pass": x_rows carries the gradient, x_all does not; dscale is the row HALF
        xr = x_rows.clone().requires_grad_(True)
        s1 = sc.clone().requires_grad_(True)
        v1 = var.clone().requires_grad_(True)
        n = x_all.shape[0]
        with torch.enable_grad():
            (self._k(kind, xr, x_all, s1, v1) * dk_rows[:, :n]).sum().backward()
        return xr.grad, (s1.grad * 0.5 if need_scale else None), (v1.grad if need_var else None)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _problem(n=150, d=6, seed=3):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(n, d, generator=g)
    y = torch.randn(n, 1, generator=g)
    sc = (0.1 * torch.randn(d, generator=g)) - 1.0
    var = torch.tensor(0.3)
    ls = torch.tensor(-2.0)
    return x, y, sc, var, ls


def _worker(rank, world, port, out, kind, logdet, cuts, block=32):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    x, y, sc, var, ls = _problem()
    if kind != "rbf":
        sc = sc[:1].clone()
    r0, r1 = cuts[rank], cuts[rank + 1]
    xl = x[r0:r1].clone().requires_grad_(True)
    yl = y[r0:r1].clone().requires_grad_(True)
    sc, var, ls = (t.clone().requires_grad_(True) for t in (sc, var, ls))
    loss = sharded_gp.sharded_exact_gp_loss(xl, yl, sc, var, ls, kind=kind, exact_logdet=(logdet == "exact"), block=block,
                                            kernels=OracleKernels())
    (3.0 * loss).sum().backward()
    torch.save({"loss": loss.detach(), "dx": xl.grad, "dy": yl.grad, "dsc": sc.grad, "dvar": var.grad, "dls": ls.grad},
               out % rank)
    dist.barrier()
    dist.destroy_process_group()


def _check(tmp_path, kind, logdet, cuts):
    _check_custom(tmp_path, kind, logdet, cuts, block=32)


def _check_custom(tmp_path, kind, logdet, cuts, block):
    world = len(cuts) - 1
    out = str(tmp_path / "r%d.pt")
    mp.spawn(_worker, args=(world, _free_port(), out, kind, logdet, cuts, block), nprocs=world, join=True)
    res = [torch.load(out % r) for r in range(world)]
    x, y, sc, var, ls = _problem()
    if kind != "rbf":
        sc = sc[:1].clone()
    x, y, sc, var, ls = (t.double().requires_grad_(True) for t in (x, y, sc, var, ls))
    want = gp_ref.exact_gp_loss(x, y, sc if kind == "rbf" else sc.reshape(()), var, ls,
                                kernel="rbf" if kind == "rbf" else "matern52_safe", logdet=logdet)
    (3.0 * want).sum().backward()

    def close(a, b, tol=2e-4):
        return float((a.double() - b.double()).norm()) <= tol * max(float(b.double().norm()), 1e-6)

    for r in range(world):
        assert res[r]["loss"].shape == (1, 1)
        assert close(res[r]["loss"], want.detach())
        assert close(res[r]["dx"], x.grad[cuts[r]:cuts[r + 1]], 2e-3)
        assert close(res[r]["dy"], y.grad[cuts[r]:cuts[r + 1]], 2e-3)
    # hyper-parameter gradients are per-rank partial sums
    assert close(sum(r_["dsc"] for r_ in res), sc.grad.reshape(res[0]["dsc"].shape), 2e-3)
    assert close(sum(r_["dvar"] for r_ in res), var.grad, 2e-3)
    assert close(sum(r_["dls"] for r_ in res), ls.grad, 2e-3)


def test_sharded_exact_gp_two_ranks_rbf_trace(tmp_path):
    _check(tmp_path, "rbf", "trace", [0, 70, 150])      # uneven row shards, 5 column blocks of 32 (last one 22)


def test_sharded_exact_gp_two_ranks_matern_exact_logdet(tmp_path):
    _check(tmp_path, "matern52", "exact", [0, 90, 150])


def test_sharded_exact_gp_three_ranks_one_without_rows_and_one_without_blocks(tmp_path):
    # 150 rows in blocks of 64 -> 3 column blocks, one per rank; rank 0 contributes no rows at all
    _check_custom(tmp_path, "rbf", "trace", [0, 0, 80, 150], block=64)
    # blocks of 96 -> 2 column blocks: rank 2 owns none of the sweep but still owns rows of the gradient
    _check_custom(tmp_path, "rbf", "exact", [0, 50, 100, 150], block=96)


def test_sharded_exact_gp_single_process_matches_oracle():
    x, y, sc, var, ls = _problem(n=97)
    xs = [t.clone().requires_grad_(True) for t in (x, y, sc, var, ls)]
    loss = sharded_gp.sharded_exact_gp_loss(*xs, kind="rbf", block=32, kernels=OracleKernels())
    loss.sum().backward()
    ref = [t.double().requires_grad_(True) for t in (x, y, sc, var, ls)]
    want = gp_ref.exact_gp_loss(*ref, kernel="rbf", logdet="trace")
    want.sum().backward()
    assert abs(float(loss) - float(want)) < 2e-4 * abs(float(want))
    for got, w in zip(xs, ref):
        assert float((got.grad.double() - w.grad.reshape(got.grad.shape)).norm()) <= 2e-3 * float(w.grad.norm())


def test_layout_of_owned_blocks():
    sw = sharded_gp.ShardedSweep(1000, 256, rank=1, world=3, group=None, dev=torch.device("cpu"), need_grad=True,
                                 kernels=OracleKernels())
    assert sw.bounds == [(0, 256), (256, 512), (512, 768), (768, 1000)]
    assert [j for j in range(sw.nJ) if sw.owned(j)] == [1]
    assert sw.rows_of(1) == (1000 - 256) + 1 + 512          # K rows 256.., y, identity rows 0..512
    assert sw.panel_rows(1) == (1000 - 512) + 1 + 512
    assert sw.panel_rows(3) == 0 + 1 + 1000
