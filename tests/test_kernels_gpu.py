"""Per-kernel parity: each C-ABI kernel against the CPU oracle on the same seeded inputs.

Tolerances: index work is bit-exact; fp32 results within 1e-4 relative of the oracle (north_star);
gradients within 1e-3 (they are compared with fp64 autograd of the oracle)."""
import pytest
import torch

import pinot_b200 as pb
from pinot_b200 import ops, synthetic
from conftest import rel_err, max_rel_err
from oracle import dgl_ref, gp_ref

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _ref_graph(g, dtype=torch.float32):
    s, d = g.edges()
    return dgl_ref.from_arrays(s.cpu(), d.cpu(), g.number_of_nodes(), g.batch_num_nodes().tolist(),
                               g.batch_num_edges().tolist())


# ---------------------------------------------------------------------------------------------
# kernel (1)
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("feat", [128, 32, 20, 7, 1])
@pytest.mark.parametrize("norm", ["both", "mean", "none"])
def test_aggregate_fwd_bwd(feat, norm):
    torch.manual_seed(feat)
    g, _ = synthetic.make_batch(40, seed=11, mean_atoms=10, std_atoms=6)
    single = pb.BatchedGraph(num_nodes=1)   # zero-in-degree node
    single.ndata["h"] = torch.zeros(1, 117)
    g = pb.batch([g, single])
    rg = _ref_graph(g)
    n = g.number_of_nodes()
    x = torch.randn(n, feat)
    go = torch.randn(n, feat)
    xr = x.double().requires_grad_(True)
    want = dgl_ref.graph_conv(rg, xr, None, None, {"both": "both", "mean": "right", "none": "none"}[norm])
    want.backward(go.double())
    gd = g.to(DEV)
    xd = x.to(DEV).requires_grad_(True)
    got = ops.aggregate(xd, gd.csr(), norm)
    got.backward(go.to(DEV))
    assert rel_err(got, want) < 1e-6
    # element-wise: sums of +-O(1) terms cancel, so the floor is 1e-3 of the largest entry
    assert max_rel_err(got, want, floor=1e-3) < 1e-4
    assert rel_err(xd.grad, xr.grad) < 1e-6


def test_aggregate_empty_graph_and_no_edges():
    g = pb.batch([pb.BatchedGraph(num_nodes=2), pb.BatchedGraph(num_nodes=1)])
    gd = g.to(DEV)
    x = torch.randn(3, 8, device=DEV)
    assert torch.count_nonzero(ops.aggregate(x, gd.csr(), "both")) == 0
    out = ops.segment_readout(x, gd.csr())
    assert torch.allclose(out[0], x[:2].sum(0)) and torch.allclose(out[1], x[2])


@pytest.mark.parametrize("feat", [128, 33])
@pytest.mark.parametrize("mean", [False, True])
def test_segment_readout_fwd_bwd(feat, mean):
    g, _ = synthetic.make_batch(50, seed=12, mean_atoms=9, std_atoms=5)
    rg = _ref_graph(g)
    n = g.number_of_nodes()
    x = torch.randn(n, feat)
    go = torch.randn(50, feat)
    xr = x.double().requires_grad_(True)
    rg.ndata["h"] = xr
    want = dgl_ref.mean_nodes(rg, "h") if mean else dgl_ref.sum_nodes(rg, "h")
    want.backward(go.double())
    gd = g.to(DEV)
    xd = x.to(DEV).requires_grad_(True)
    gd.ndata["h"] = xd
    got = (pb.graph.mean_nodes if mean else pb.graph.sum_nodes)(gd, "h")
    got.backward(go.to(DEV))
    assert got.shape == (50, feat)
    assert rel_err(got, want) < 1e-6 and rel_err(xd.grad, xr.grad) < 1e-6


# ---------------------------------------------------------------------------------------------
# kernel (2)
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("n,fin,fout", [(1000, 117, 128), (517, 128, 128), (300, 128, 32), (131, 20, 7), (1, 5, 3)])
@pytest.mark.parametrize("w_is_out_in", [True, False])
@pytest.mark.parametrize("act", [None, "tanh"])
def test_node_update_fwd_bwd(n, fin, fout, w_is_out_in, act):
    torch.manual_seed(n + fin)
    x = torch.randn(n, fin)
    w = torch.randn(fout, fin) * 0.1 if w_is_out_in else torch.randn(fin, fout) * 0.1
    b = torch.randn(fout)
    go = torch.randn(n, fout)
    xr, wr, br = (t.double().requires_grad_(True) for t in (x, w, b))
    want = xr @ (wr.t() if w_is_out_in else wr) + br
    if act:
        want = torch.tanh(want)
    want.backward(go.double())
    xd, wd, bd = (t.to(DEV).requires_grad_(True) for t in (x, w, b))
    got = ops.node_update(xd, wd, bd, w_is_out_in=w_is_out_in, act=act)
    got.backward(go.to(DEV))
    assert rel_err(got, want) < 2e-6
    assert rel_err(xd.grad, xr.grad) < 1e-5
    assert rel_err(wd.grad, wr.grad) < 1e-5
    assert rel_err(bd.grad, br.grad) < 1e-5


def test_node_update_large_rows_uses_split_k_weight_gradient():
    torch.manual_seed(0)
    n, f = 30000, 128
    x = torch.randn(n, f, device=DEV)
    w = (0.05 * torch.randn(f, f, device=DEV)).requires_grad_(True)
    b = torch.zeros(f, device=DEV, requires_grad=True)
    go = torch.randn(n, f, device=DEV)
    y = ops.node_update(x, w, b, w_is_out_in=False, act="tanh")
    y.backward(go)
    yr = torch.tanh(x.double() @ w.detach().double())
    dz = go.double() * (1 - yr * yr)
    assert rel_err(y, yr) < 2e-6
    assert rel_err(w.grad, x.double().t() @ dz) < 1e-5
    assert rel_err(b.grad, dz.sum(0)) < 1e-5


@pytest.mark.parametrize("n,fin,fout", [(20000, 128, 128), (9001, 128, 32), (8200, 64, 128), (12345, 36, 20)])
@pytest.mark.parametrize("w_is_out_in", [True, False])
@pytest.mark.parametrize("act", [None, "tanh"])
def test_node_update_tall_inputs_take_the_persistent_tma_kernel(n, fin, fout, w_is_out_in, act):
    """n >= 8192 rows with at most 128 features go through gemm_node.cu (forward and dX); same answers as
    fp64, and the same as the register-staged kernel it replaces (PINOT_B200_NODE_GEMM=0 is read once per
    process, so that comparison is on a row count below the switch-over)."""
    torch.manual_seed(n + fin)
    x = torch.randn(n, fin)
    w = torch.randn(fout, fin) * 0.1 if w_is_out_in else torch.randn(fin, fout) * 0.1
    b = torch.randn(fout)
    go = torch.randn(n, fout)
    xd, wd, bd = (t.to(DEV).requires_grad_(True) for t in (x, w, b))
    got = ops.node_update(xd, wd, bd, w_is_out_in=w_is_out_in, act=act)
    got.backward(go.to(DEV))
    wm = (w.t() if w_is_out_in else w).double()
    want = x.double() @ wm + b.double()
    dz = go.double()
    if act:
        want = torch.tanh(want)
        dz = dz * (1 - want * want)
    assert rel_err(got, want) < 2e-6
    assert rel_err(xd.grad, dz @ wm.t()) < 1e-5
    assert rel_err(bd.grad, dz.sum(0)) < 1e-5
    # rows below the switch-over run on the other kernel: the shared rows must agree to the last bits
    small = ops.node_update(xd[:4000].detach(), wd.detach(), bd.detach(), w_is_out_in=w_is_out_in, act=act)
    assert rel_err(got[:4000], small) < 1e-6


def test_node_update_tall_accumulate_and_ragged_tail():
    torch.manual_seed(3)
    n = 8192 + 77                                   # last row tile is partial
    x, hn = torch.randn(n, 128, device=DEV), torch.randn(n, 128, device=DEV)
    ws, wn = torch.randn(64, 128, device=DEV) * 0.1, torch.randn(64, 128, device=DEV) * 0.1
    bs, bn = torch.randn(64, device=DEV), torch.randn(64, device=DEV)
    y = ops.node_update(x, ws, bs, w_is_out_in=True)
    y = ops.node_update(hn, wn, bn, y_prev=y, w_is_out_in=True, act="tanh")
    want = torch.tanh(x.double() @ ws.double().t() + bs.double() + hn.double() @ wn.double().t() + bn.double())
    assert rel_err(y, want) < 2e-6


def test_node_update_accumulate_for_sage():
    torch.manual_seed(1)
    x, hn = torch.randn(200, 64, device=DEV), torch.randn(200, 64, device=DEV)
    ws, wn = torch.randn(48, 64, device=DEV) * 0.1, torch.randn(48, 64, device=DEV) * 0.1
    bs, bn = torch.randn(48, device=DEV), torch.randn(48, device=DEV)
    y = ops.node_update(x, ws, bs, w_is_out_in=True)
    y = ops.node_update(hn, wn, bn, y_prev=y, w_is_out_in=True, act="tanh")
    want = torch.tanh(x.double() @ ws.double().t() + bs.double() + hn.double() @ wn.double().t() + bn.double())
    assert rel_err(y, want) < 2e-6


@pytest.mark.parametrize("ta", [False, True])
@pytest.mark.parametrize("tb", [False, True])
@pytest.mark.parametrize("m,n,k", [(257, 130, 77), (64, 64, 4096), (5, 3, 2), (128, 128, 128), (300, 1, 50)])
def test_sgemm_all_layouts(ta, tb, m, n, k):
    torch.manual_seed(m * n + k)
    a = torch.randn((k, m) if ta else (m, k))
    b = torch.randn((n, k) if tb else (k, n))
    want = (a.double().t() if ta else a.double()) @ (b.double().t() if tb else b.double())
    ad, bd = a.to(DEV).requires_grad_(True), b.to(DEV).requires_grad_(True)
    got = ops.matmul(ad, bd, ta, tb)
    assert rel_err(got, want) < 2e-6
    go = torch.randn(m, n)
    got.backward(go.to(DEV))
    ar, br = a.double().requires_grad_(True), b.double().requires_grad_(True)
    ((ar.t() if ta else ar) @ (br.t() if tb else br)).backward(go.double())
    assert rel_err(ad.grad, ar.grad) < 1e-5 and rel_err(bd.grad, br.grad) < 1e-5


# ---------------------------------------------------------------------------------------------
# kernel (3)
# ---------------------------------------------------------------------------------------------
def _kernel_params(kind, d, scale, variance):
    sc = torch.full((d,), scale) if kind == "rbf" else torch.tensor(scale)
    return sc, torch.tensor(variance)


@pytest.mark.parametrize("m,n,k", [(300, 200, 100), (1024, 1152, 1030), (130, 72, 36), (128, 128, 2048)])
def test_sgemm_nt_tma_matches_fp64_and_register_kernel(m, n, k):
    """TMA-fed kernel (pre-split operands) = the same 3xTF32 arithmetic as pinot_b200_sgemm(0, 1, ...)."""
    from pinot_b200 import _lib
    from pinot_b200._lib import call, ptr
    g = torch.Generator().manual_seed(m + n + k)
    lda = ops._pad4(k)
    a = torch.zeros(m, lda)
    b = torch.zeros(n, lda)
    a[:, :k] = torch.randn(m, k, generator=g)
    b[:, :k] = torch.randn(n, k, generator=g)
    c0 = torch.randn(m, n, generator=g)
    want = 0.5 * (a[:, :k].double() @ b[:, :k].double().t()) - 2.0 * c0.double()
    a, b = a.to(DEV), b.to(DEV)
    nbytes = int(_lib.load().pinot_b200_sgemm_nt_tma_workspace_bytes(m, n, k))
    ws = torch.empty(nbytes // 4, device=DEV)
    got = c0.clone().to(DEV)
    call("sgemm_nt_tma", m, n, k, 0.5, ptr(a), lda, ptr(b), lda, -2.0, ptr(got), n, ptr(ws), nbytes)
    reg = c0.clone().to(DEV)
    call("sgemm", 0, 1, m, n, k, 0.5, ptr(a), lda, ptr(b), lda, -2.0, ptr(reg), n)
    assert rel_err(got, want) < 1e-5
    assert rel_err(got, reg) < 1e-6
    with pytest.raises(_lib.PinotB200Error, match="workspace too small"):
        call("sgemm_nt_tma", m, n, k, 1.0, ptr(a), lda, ptr(b), lda, 0.0, ptr(got), n, ptr(ws), nbytes - 4)


@pytest.mark.parametrize("kind", ["rbf", "matern52"])
@pytest.mark.parametrize("n,n2,d", [(300, 300, 128), (131, 67, 32), (64, 200, 7), (1, 5, 3), (70, 70, 200)])
def test_gram_forward(kind, n, n2, d):
    torch.manual_seed(n + d)
    x, y = torch.randn(n, d) * 0.6, torch.randn(n2, d) * 0.6
    y[: min(n, n2) // 2] = x[: min(n, n2) // 2]      # duplicate rows: exact zero distances off the diagonal too
    sc, var = _kernel_params(kind, d, -0.7, 0.4)
    want = gp_ref.KERNELS[kind](x, y, sc, var)
    got = ops.gram(kind, x.to(DEV), y.to(DEV), sc.to(DEV), var.to(DEV))
    assert got.shape == (n, n2)
    assert max_rel_err(got, want) < 1e-4, "K must match the difference-form reference to 1e-4 relative"
    # fused _perturb on a rectangular matrix: +eps on the main diagonal only
    got_p = ops.gram(kind, x.to(DEV), y.to(DEV), sc.to(DEV), var.to(DEV), diag_add=1e-5)
    assert max_rel_err(got_p, gp_ref.perturb(want)) < 1e-4
    # K(x, x): exact amplitude on the diagonal
    kxx = ops.gram(kind, x.to(DEV), None, sc.to(DEV), var.to(DEV))
    assert torch.allclose(torch.diagonal(kxx).cpu(), torch.exp(var).expand(n), rtol=1e-6)
    assert torch.equal(kxx, kxx.t())


@pytest.mark.parametrize("kind", ["rbf", "matern52"])
@pytest.mark.parametrize("n,n2,d", [(150, 150, 128), (97, 41, 32), (40, 90, 7), (33, 33, 200)])
def test_gram_backward_two_argument(kind, n, n2, d):
    torch.manual_seed(n * 3 + d)
    x, y = torch.randn(n, d) * 0.6, torch.randn(n2, d) * 0.6
    y[:5] = x[:5]
    sc, var = _kernel_params(kind, d, -0.5, 0.2)
    gk = torch.randn(n, n2)
    ref_kind = "matern52_safe" if kind == "matern52" else kind
    xr, yr, scr, varr = (t.double().requires_grad_(True) for t in (x, y, sc, var))
    gp_ref.KERNELS[ref_kind](xr, yr, scr, varr).backward(gk.double())
    xd, yd, scd, vard = (t.to(DEV).requires_grad_(True) for t in (x, y, sc, var))
    ops.gram(kind, xd, yd, scd, vard).backward(gk.to(DEV))
    assert rel_err(xd.grad, xr.grad) < 1e-3
    assert rel_err(yd.grad, yr.grad) < 1e-3
    assert rel_err(scd.grad, scr.grad) < 1e-3
    assert rel_err(vard.grad, varr.grad) < 1e-3


@pytest.mark.parametrize("kind", ["rbf", "matern52"])
def test_gram_backward_same_argument(kind):
    torch.manual_seed(5)
    n, d = 140, 64
    x = torch.randn(n, d) * 0.6
    x[7] = x[3]
    sc, var = _kernel_params(kind, d, -0.5, 0.2)
    gk = torch.randn(n, n)
    ref_kind = "matern52_safe" if kind == "matern52" else kind
    xr, scr, varr = (t.double().requires_grad_(True) for t in (x, sc, var))
    gp_ref.KERNELS[ref_kind](xr, xr, scr, varr).backward(gk.double())
    xd, scd, vard = (t.to(DEV).requires_grad_(True) for t in (x, sc, var))
    ops.gram(kind, xd, None, scd, vard).backward(gk.to(DEV))
    assert rel_err(xd.grad, xr.grad) < 1e-3
    assert rel_err(scd.grad, scr.grad) < 1e-3 and rel_err(vard.grad, varr.grad) < 1e-3
    assert torch.isfinite(xd.grad).all()


def _embedding_like(n, d, seed, offset=3.0, spread=1.0):
    """Rows with a large common component (like sum-pooled tanh embeddings: |x| >> typical distance), exact
    duplicates and near-duplicates -- the inputs on which the GEMM form of the distance loses its digits."""
    g = torch.Generator().manual_seed(seed)
    x = offset + spread * torch.randn(n, d, generator=g)
    x[5] = x[2]                                   # exact duplicate
    x[9] = x[4] + 1e-4 * torch.randn(d, generator=g)   # near duplicates at several scales
    x[11] = x[4] + 1e-2 * torch.randn(d, generator=g)
    x[17] = x[6] + 1e-6
    return x


@pytest.mark.parametrize("kind", ["rbf", "matern52"])
@pytest.mark.parametrize("n,n2,d,scale", [(700, 700, 128, -3.0), (1128, 300, 128, -1.0), (300, 1500, 32, -2.0),
                                          (513, 513, 100, 0.0), (260, 257, 8, -1.0), (2304, 2304, 128, -3.0)])
def test_gram_forward_tensor_core_route(kind, n, n2, d, scale, monkeypatch):
    """The tcgen05 route (csrc/gram_tc.cu) on small and ragged shapes (forced: by default it starts at ~2000^2
    pairs, below which the CUDA-core kernel is faster; (2304, 2304) takes it unforced).  Against the fp64 difference-form reference, 1e-4
    element-wise (entries below 1e-6 of the largest measured against that floor), including exact and near
    duplicates where |x|^2 + |y|^2 - 2 x.y cancels, and against the CUDA-core route."""
    from pinot_b200 import _lib
    from pinot_b200._lib import call, ptr
    if n * n2 < 4000000:
        monkeypatch.setenv("PINOT_B200_GRAM", "tc")     # read by the library on every call
    x = _embedding_like(n, d, n + d)
    y = x if n == n2 else _embedding_like(n2, d, n2 + d + 1)
    if n != n2:
        y[:40] = x[:40]                            # duplicates across the two arguments
    sc, var = _kernel_params(kind, d, scale, 0.4)
    want = gp_ref.KERNELS[kind](x.double(), y.double(), sc.double(), var.double())
    xd, yd = x.to(DEV), (x.to(DEV) if n == n2 else y.to(DEV))
    if n == n2:
        yd = xd
    assert _lib.gram_workspace_bytes(n, n2, d, n == n2, False) > 0
    got = ops.gram_raw(kind, xd, yd, sc.to(DEV), var.to(DEV).reshape(1), diag_add=1e-5)
    want_p = gp_ref.perturb(want)
    assert max_rel_err(got, want_p) < 1e-4
    assert rel_err(got, want_p) < 1e-6
    # the CUDA-core route (no workspace) on the same inputs
    ffma = torch.empty((n, n2), device=DEV)
    scd, vard = sc.to(DEV).reshape(-1).contiguous(), var.to(DEV).reshape(1)     # kept alive across the launch
    call("gram_fwd", ops.KIND[kind], ptr(xd), ptr(yd), ptr(scd), scd.numel(), ptr(vard), ptr(ffma), n2, n, n2, d, 1e-5, None, 0,
         None, 0)
    assert max_rel_err(got, ffma) < 1e-4
    if n == n2:
        assert torch.allclose(torch.diagonal(got).cpu(), (torch.exp(var) + 1e-5).expand(n), rtol=1e-6)
        # lower_only writes whole tiles on or below the diagonal and leaves the rest alone
        lo = torch.full((n, n), -7.0, device=DEV)
        ops.gram_raw(kind, xd, xd, sc.to(DEV), var.to(DEV).reshape(1), out=lo, ldo=n, diag_add=1e-5, lower_only=True)
        assert torch.equal(torch.tril(lo), torch.tril(got))
        assert float(lo[0, n - 1]) == -7.0


@pytest.mark.parametrize("kind", ["rbf", "matern52"])
@pytest.mark.parametrize("n,n2,d,scale,same", [(600, 600, 128, -3.0, True), (400, 900, 64, -2.0, False),
                                               (300, 300, 36, -1.0, False), (1500, 1500, 128, -3.0, True)])
def test_gram_backward_tensor_core_route(kind, n, n2, d, scale, same, monkeypatch):
    """Backward through the coefficient-matrix GEMMs (csrc/gram_tc.cu) against fp64 autograd of the reference
    formula; the near pairs go through the exact fix-up, the others through rowsum(c) xs - c ys.  Forced onto
    the tensor-core route (by default it starts at ~1200^2 pairs; (1500, 1500) takes it unforced)."""
    if n * n2 < 1500000:
        monkeypatch.setenv("PINOT_B200_GRAM", "tc")
    from pinot_b200 import _lib
    assert _lib.gram_workspace_bytes(n, n if same else n2, d, same, True, not same) > 0
    x = _embedding_like(n, d, 3 * n + d, offset=2.0)
    y = None if same else _embedding_like(n2, d, 5 * n2 + d, offset=2.0)
    if not same:
        y[:30] = x[:30]
    sc, var = _kernel_params(kind, d, scale, 0.2)
    g = torch.Generator().manual_seed(1)
    gk = torch.randn(n, n if same else n2, generator=g)
    if same:
        gk = 0.5 * (gk + gk.t())
    ref_kind = "matern52_safe" if kind == "matern52" else kind
    xr, scr, varr = (t.double().requires_grad_(True) for t in (x, sc, var))
    yr = xr if same else y.double().requires_grad_(True)
    gp_ref.KERNELS[ref_kind](xr, yr, scr, varr).backward(gk.double())
    xd, scd, vard = (t.to(DEV).requires_grad_(True) for t in (x, sc, var))
    yd = None if same else y.to(DEV).requires_grad_(True)
    ops.gram(kind, xd, yd, scd, vard).backward(gk.to(DEV))
    # rows whose nearest neighbour sits 1e-6 ... 1e-4 away: fp32 inputs cannot resolve that difference to 1e-3 (the
    # reference's own fp32 arithmetic cannot either); they are held to 2e-2, everything else to 1e-3
    close = [4, 6, 9, 11, 17]
    far = [i for i in range(n) if i not in close]
    assert rel_err(xd.grad[far], xr.grad[far]) < 1e-3
    assert rel_err(xd.grad[close], xr.grad[close]) < 2e-2
    if not same:
        assert rel_err(yd.grad, yr.grad) < 5e-3
        assert rel_err(yd.grad[30:], yr.grad[30:]) < 1e-3
    assert rel_err(scd.grad, scr.grad) < 5e-3
    assert rel_err(vard.grad, varr.grad) < 1e-3
    assert torch.isfinite(xd.grad).all()
    # the symmetric single-pass mode used by the exact-GP loss (row pass, doubled by the caller)
    if same:
        dx, _, dscale, dvar = ops.gram_bwd_raw(kind, xd.detach(), xd.detach(), scd.detach().reshape(-1),
                                               vard.detach().reshape(1), gk.to(DEV), need_y=False, symmetric=True)
        assert rel_err(2.0 * dx[far], xr.grad[far]) < 1e-3
        assert rel_err(2.0 * dscale.reshape(scr.shape), scr.grad) < 5e-3
        assert rel_err(dvar.reshape(()), varr.grad) < 1e-3


# ---------------------------------------------------------------------------------------------
# kernel (4)
# ---------------------------------------------------------------------------------------------
def _spd(n, seed, cond_shift=0.5):
    g = torch.Generator().manual_seed(seed)
    a = torch.randn(n, n, generator=g, dtype=torch.float64)
    return (a @ a.t() / n + cond_shift * torch.eye(n, dtype=torch.float64))


@pytest.mark.parametrize("n,r", [(64, 0), (65, 3), (200, 17), (513, 300), (1000, 1), (1, 1), (130, 0)])
def test_potrf_ext_sweep(n, r):
    k = _spd(n, n)
    b = torch.randn(r, n, dtype=torch.float64)
    ld = ops._pad4(n)
    e = torch.zeros(n + r, ld, device=DEV)
    e[:n, :n] = k.float().to(DEV)
    e[n:, :n] = b.float().to(DEV)
    diag, info = ops.potrf_ext_raw(e, n, n_ext_dense=r)
    assert int(info.item()) == 0
    l_want = torch.linalg.cholesky(k)
    l_got = torch.tril(e[:n, :n]).cpu()
    assert rel_err(l_got, l_want) < 1e-5
    assert rel_err(diag, torch.diagonal(l_want)) < 1e-5
    if r:
        x_want = torch.linalg.solve_triangular(l_want, b.t(), upper=False).t()   # B L^-T
        assert rel_err(e[n:, :n], x_want) < 1e-4


@pytest.mark.parametrize("n", [64, 100, 300])
def test_potrf_ext_identity_rows_give_inverse_transpose(n):
    k = _spd(n, 7 * n)
    y = torch.randn(n, dtype=torch.float64)
    ld = ops._pad4(n)
    e = torch.zeros(2 * n + 1, ld, device=DEV)
    e[:n, :n] = k.float().to(DEV)
    e[n, :n] = y.float().to(DEV)
    e[n + 1:, :n].diagonal().fill_(1.0)
    ops.potrf_ext_raw(e, n, n_ext_dense=1, ext_identity=True)
    l = torch.linalg.cholesky(k)
    linv_t = torch.linalg.inv(l).t()
    got = e[n + 1:, :n].cpu()
    assert rel_err(got, linv_t) < 1e-4
    assert torch.count_nonzero(torch.tril(got, -1)) == 0, "L^-T must stay exactly upper triangular"
    assert rel_err(e[n, :n], torch.linalg.solve_triangular(l, y[:, None], upper=False).flatten()) < 1e-4


def test_potrf_two_level_sweep_with_and_without_tma_scratch():
    """n >= 6144 runs rank-512 outer updates; given the scratch tail they go through the TMA-fed kernel.
    Both routes must give the same factor, and the factor must reproduce the matrix."""
    from pinot_b200 import _lib
    from pinot_b200._lib import call, ptr
    n, r = 6400, 3
    g = torch.Generator().manual_seed(5)
    a = torch.randn(n, 64, generator=g).to(DEV)
    k = a @ a.t() / 64 + 2.0 * torch.eye(n, device=DEV)
    ld = ops._pad4(n)
    e0 = torch.zeros(2 * n + r, ld, device=DEV)
    e0[:n, :n] = k
    e0[n:n + r, :n] = torch.randn(r, n, generator=g).to(DEV)
    e0[n + r:, :n].diagonal().fill_(1.0)
    outs = []
    for with_scratch in (False, True):
        nbytes = _lib.potrf_workspace_bytes(n, r) if with_scratch else _lib.potrf_workspace_bytes(n)
        ws = torch.empty(nbytes // 4, device=DEV)
        e = e0.clone()
        info = torch.zeros(1, dtype=torch.int32, device=DEV)
        call("potrf_ext", ptr(e), ld, n, r, 1, None, ptr(ws), nbytes, ptr(info))
        assert int(info.item()) == 0
        outs.append(e)
    assert _lib.potrf_workspace_bytes(n, r) > _lib.potrf_workspace_bytes(n)
    assert rel_err(outs[1], outs[0]) < 1e-6
    l = torch.tril(outs[1][:n, :n]).double()
    assert rel_err(l @ l.t(), k.double()) < 1e-5
    w = outs[1][n + r:, :n].double()                       # L^-T
    assert rel_err(l.t() @ w, torch.eye(n, dtype=torch.float64, device=DEV)) < 1e-4


@pytest.mark.parametrize("n,r,ident", [(1000, 3, 1), (4200, 1, 1), (1128, 1, 1), (129, 0, 1), (64, 2, 0), (777, 900, 0),
                                       (2050, 5, 1), (5000, 1, 0)])
def test_single_launch_sweep_matches_launch_per_panel_sweep(n, r, ident):
    """PINOT_B200_SWEEP=persistent (and a workspace with the flag tail) selects the persistent cooperative kernel
    (csrc/chol_sweep.cu) below n = 6144; otherwise the sweep runs one launch per panel.  Same arithmetic,
    different order of the fp32 additions: they must agree to rounding, and the factor must reproduce the matrix."""
    import os
    from pinot_b200 import _lib
    from pinot_b200._lib import call, ptr
    g = torch.Generator().manual_seed(n + r)
    a = torch.randn(n, 96, generator=g).to(DEV)
    k = a @ a.t() / 96 + 1.5 * torch.eye(n, device=DEV)
    ld = ops._pad4(n)
    rows = n + r + (n if ident else 0)
    e0 = torch.zeros(rows, ld, device=DEV)
    e0[:n, :n] = torch.tril(k) + 123.0 * torch.triu(torch.ones(n, n, device=DEV), 1)   # junk above the diagonal
    e0[n:n + r, :n] = torch.randn(r, n, generator=g).to(DEV)
    if ident:
        e0[n + r:, :n].diagonal().fill_(1.0)
    outs, diags = [], []
    assert _lib.potrf_workspace_bytes(n, r) > _lib.potrf_workspace_bytes(n)
    for tail in (False, True):
        nbytes = _lib.potrf_workspace_bytes(n, r)
        ws = torch.empty(nbytes // 4, device=DEV)
        e = e0.clone()
        info = torch.full((1,), -3, dtype=torch.int32, device=DEV)
        diag = torch.empty(n, device=DEV)
        saved = os.environ.get("PINOT_B200_SWEEP")
        os.environ["PINOT_B200_SWEEP"] = "persistent" if tail else "launch"
        try:
            for _ in range(2):                      # twice: stale flags in the workspace must not matter
                e.copy_(e0)
                call("potrf_ext", ptr(e), ld, n, r, ident, ptr(diag), ptr(ws), nbytes, ptr(info))
        finally:
            if saved is None:
                os.environ.pop("PINOT_B200_SWEEP", None)
            else:
                os.environ["PINOT_B200_SWEEP"] = saved
        assert int(info.item()) == 0
        outs.append(e)
        diags.append(diag)
    lo, hi = outs
    assert rel_err(torch.tril(hi[:n, :n]), torch.tril(lo[:n, :n])) < 2e-6
    assert rel_err(diags[1], diags[0]) < 2e-6
    assert torch.equal(torch.triu(hi[:n, :n], 1), torch.triu(e0[:n, :n], 1)), "the strict upper triangle is not touched"
    if r:
        assert rel_err(hi[n:n + r, :n], lo[n:n + r, :n]) < 1e-5
    if ident:
        assert rel_err(hi[n + r:, :n], lo[n + r:, :n]) < 1e-5
        assert torch.count_nonzero(torch.tril(hi[n + r:, :n], -1)) == 0
    l = torch.tril(hi[:n, :n]).double()
    assert rel_err(l @ l.t(), k.double()) < 1e-5


def test_gp_nll_grad_tma_workspace_route_matches_register_route():
    from pinot_b200 import _lib
    from pinot_b200._lib import call, ptr
    n = 1300
    g = torch.Generator().manual_seed(11)
    w = (torch.triu(torch.randn(n, n, generator=g)) * 0.05).to(DEV)
    diag = (torch.rand(n, generator=g) + 0.5).to(DEV)
    alpha = torch.randn(n, generator=g).to(DEV)
    up = torch.full((1,), 0.7, device=DEV)
    for exact in (0, 1):
        wd = w.double() * (1.0 if exact else diag.double())
        want = 0.7 * (0.5 * wd @ w.double().t() - 0.5 * torch.outer(alpha.double(), alpha.double()))
        res = []
        for nbytes in (0, _lib.gp_nll_grad_workspace_bytes(n)):
            ws = torch.empty(nbytes // 4, device=DEV) if nbytes else None
            dk = torch.full((n, n), float("nan"), device=DEV)
            call("gp_nll_grad", ptr(w), n, n, ptr(diag), ptr(alpha), ptr(up), exact, ptr(dk), n, ptr(ws), nbytes)
            res.append(dk)
        assert _lib.gp_nll_grad_workspace_bytes(n) > 0
        assert rel_err(res[0], want) < 1e-5 and rel_err(res[1], want) < 1e-5
        assert rel_err(res[1], res[0]) < 1e-6
        assert torch.equal(res[1], res[1].t())              # mirrored: exactly symmetric


def test_potrf_reports_non_positive_definite_like_torch():
    n = 150
    k = _spd(n, 3).float()
    k[100, 100] = -1.0
    e = k.to(DEV).clone()
    with pytest.raises(RuntimeError, match="not positive-definite"):
        ops.potrf_ext_raw(e, n)
    e = k.to(DEV).clone()
    _, info = ops.potrf_ext_raw(e, n, check=False)
    assert int(info.item()) == 101      # LAPACK convention: order of the failing leading minor


@pytest.mark.parametrize("n,r", [(70, 5), (300, 64)])
def test_trsm_rlt(n, r):
    l = torch.linalg.cholesky(_spd(n, n + 1))
    b = torch.randn(r, n, dtype=torch.float64)
    ld, xd = l.float().to(DEV).contiguous(), b.float().to(DEV).contiguous()
    pb._lib.call("trsm_rlt", ld.data_ptr(), n, n, xd.data_ptr(), n, r)
    want = torch.linalg.solve_triangular(l, b.t(), upper=False).t()
    assert rel_err(xd, want) < 1e-4


@pytest.mark.parametrize("n,r", [(1024, 1500), (1792, 1100), (2400, 1024)])
def test_trsm_rlt_with_workspace_matches_the_plain_solve(n, r):
    """pinot_b200_trsm_rlt_ws: the wide updates on the TMA-fed GEMM with accumulators flushed into registers every
    128 K-columns.  Against the fp64 solve, and against the plain schedule on the same inputs (both 128-wide
    accumulation: they agree to fp32 rounding); ragged last block (n = 1792 = 3.5 x 512, n = 2400)."""
    from pinot_b200 import _lib
    l = torch.linalg.cholesky(_spd(n, n + 1))
    b = torch.randn(r, n, dtype=torch.float64)
    ld = l.float().to(DEV).contiguous()
    x1, x2 = b.float().to(DEV).contiguous(), b.float().to(DEV).contiguous()
    nbytes = _lib.trsm_rlt_workspace_bytes(n, r)
    assert nbytes > 0
    ws = torch.empty(nbytes // 4, device=DEV)
    pb._lib.call("trsm_rlt_ws", ld.data_ptr(), n, n, x1.data_ptr(), n, r, ws.data_ptr(), nbytes)
    pb._lib.call("trsm_rlt", ld.data_ptr(), n, n, x2.data_ptr(), n, r)
    want = torch.linalg.solve_triangular(l, b.t(), upper=False).t()
    assert rel_err(x1, want) < 1e-4 and rel_err(x2, want) < 1e-4
    assert rel_err(x1, x2) < 2e-6
    # no workspace / too small a problem: the plain schedule, same answer
    x3 = b.float().to(DEV).contiguous()
    pb._lib.call("trsm_rlt_ws", ld.data_ptr(), n, n, x3.data_ptr(), n, r, None, 0)
    assert torch.equal(x3, x2)
    assert _lib.trsm_rlt_workspace_bytes(512, 4096) == 0


@pytest.mark.parametrize("n,r", [(90, 4), (200, 33)])
def test_chol_solve_autograd(n, r):
    torch.manual_seed(n)
    k = _spd(n, n + 5)
    b = torch.randn(n, r, dtype=torch.float64)
    gl = torch.tril(torch.randn(n, n, dtype=torch.float64))
    gx = torch.randn(n, r, dtype=torch.float64)
    kr, br = k.clone().requires_grad_(True), b.clone().requires_grad_(True)
    lr = torch.linalg.cholesky(kr)
    xr = torch.linalg.solve_triangular(lr, br, upper=False)
    ((lr * gl).sum() + (xr * gx).sum()).backward()
    kd, bd = k.float().to(DEV).requires_grad_(True), b.float().to(DEV).requires_grad_(True)
    l, x = ops.chol_solve(kd, bd)
    assert rel_err(l, lr) < 1e-5 and rel_err(x, xr) < 1e-4
    assert torch.count_nonzero(torch.triu(l, 1)) == 0
    ((l * gl.float().to(DEV)).sum() + (x * gx.float().to(DEV)).sum()).backward()
    kg = 0.5 * (kr.grad + kr.grad.t())
    assert rel_err(kd.grad, kg) < 2e-3
    assert rel_err(bd.grad, br.grad) < 1e-3


@pytest.mark.parametrize("kind", ["rbf", "matern52"])
@pytest.mark.parametrize("logdet", ["trace", "exact"])
@pytest.mark.parametrize("n,d", [(257, 32), (64, 128), (700, 128)])
def test_exact_gp_loss_and_gradients(kind, logdet, n, d):
    torch.manual_seed(n + d)
    x = torch.randn(n, d) * 0.5
    y = torch.randn(n, 1)
    sc, var = _kernel_params(kind, d, -0.5, 0.3)
    ls = torch.tensor(-2.0)
    ref_kind = "matern52_safe" if kind == "matern52" else kind
    xr, yr, scr, varr, lsr = (t.double().requires_grad_(True) for t in (x, y, sc, var, ls))
    want = gp_ref.exact_gp_loss(xr, yr, scr, varr, lsr, kernel=ref_kind, logdet=logdet)
    want.sum().backward()
    want32 = gp_ref.exact_gp_loss(x, y, sc, var, ls, kernel=kind, logdet=logdet)
    xd, yd, scd, vard, lsd = (t.to(DEV).requires_grad_(True) for t in (x, y, sc, var, ls))
    got = ops.exact_gp_loss(xd, yd, scd, vard, lsd, kind=kind, exact_logdet=(logdet == "exact"))
    assert got.shape == (1, 1)
    assert rel_err(got, want) < 1e-5, "NLL within 1e-3 absolute is the bar; fp32 reaches 1e-5 relative"
    assert abs(float(got) - float(want32)) < 1e-3 * max(1.0, abs(float(want32)))
    (3.0 * got.sum()).backward()
    for a, b_, name in [(xd, xr, "x"), (yd, yr, "y"), (scd, scr, "scale"), (vard, varr, "variance"), (lsd, lsr, "log_sigma")]:
        assert rel_err(a.grad, 3.0 * b_.grad) < 2e-3, name


@pytest.mark.parametrize("kind", ["rbf", "matern52"])
@pytest.mark.parametrize("n,nte,d", [(300, 50, 32), (129, 7, 128), (64, 1, 16)])
def test_exact_gp_posterior(kind, n, nte, d):
    torch.manual_seed(n + nte)
    x, xt = torch.randn(n, d) * 0.5, torch.randn(nte, d) * 0.5
    xt[0] = x[0]   # a test point equal to a training point
    y = torch.randn(n, 1)
    sc, var = _kernel_params(kind, d, -0.5, 0.3)
    ls = torch.tensor(-2.0)
    mean_w, cov_w = gp_ref.exact_gp_condition(xt.double(), x.double(), y.double(), sc.double(), var.double(),
                                              ls.double(), kernel=kind)
    args = [t.to(DEV) for t in (xt, x, y, sc, var, ls)]
    mean, cov = ops.exact_gp_posterior(*args, kind=kind, full_cov=True)
    assert rel_err(mean, mean_w) < 1e-4 and rel_err(cov, cov_w) < 1e-4
    mean2, v = ops.exact_gp_posterior(*args, kind=kind, full_cov=False)
    assert rel_err(mean2, mean_w) < 1e-4
    assert max_rel_err(v, torch.diagonal(cov_w), floor=1e-3) < 1e-3


def test_deferred_positive_definite_check():
    """check_positive_definite="deferred": loss() does not synchronise; verify() raises afterwards."""
    torch.manual_seed(0)
    reg = pb.regressors.ExactGaussianProcessRegressor(8).to(DEV)
    x = torch.randn(40, 8, device=DEV)
    x[3, 2] = float("nan")                        # poisons row/column 3 of K: pivot 3 is NaN
    y = torch.randn(40, 1, device=DEV)
    reg.check_positive_definite = True
    with pytest.raises(RuntimeError, match="not positive-definite"):
        reg.loss(x, y)
    reg.check_positive_definite = "deferred"
    reg.loss(x, y)                                # no exception here
    with pytest.raises(RuntimeError, match="not positive-definite"):
        reg.verify()
    reg.check_positive_definite = False
    reg.loss(x, y)
    reg.verify()                                  # checking switched off
