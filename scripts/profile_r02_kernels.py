"""Launches the kernels round 2 is about, a few times each, for `ncu --set full` (scripts/gpu_run_ncu_r02.sh):
  * gram_tma_kernel<1, 0> / <2, 0> / <3, 0>: Gram forward, backward coefficient pass and norm pass on the tensor
    cores at N = 4200 (the bench's embeddings: real near pairs) and forward at N = 16384 (random rows),
  * one whole Cholesky sweep at N = 4200 with the identity rows (chol_diag_kernel + 65 x tgemm_kernel<0,1,0,1,128>,
    the look-ahead update with the fused factor and panel solve),
  * trsm_rlt of 32768 test rows against a 4096 factor (the posterior's solve)."""
import os
import sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pinot_b200 as pb
from pinot_b200 import _lib, ops, synthetic
from pinot_b200._lib import call, ptr

DEV = "cuda"
which = sys.argv[1] if len(sys.argv) > 1 else "all"
torch.manual_seed(0)
if which in ("all", "gram"):
    g, y = synthetic.make_batch(4200, seed=2666)
    rep = pb.representation.Sequential(pb.representation.dgl_legacy.gn("GraphConv"), [128, "tanh", 128, "tanh", 128, "tanh"]).to(DEV)
    with torch.no_grad():
        x = rep(g.to(DEV)).contiguous()
    for n, rows in ((4200, x), (16384, (3.0 + torch.randn(16384, 128)).to(DEV))):
        d = 128
        sc = torch.full((d,), -3.0, device=DEV)
        var = torch.zeros(1, device=DEV)
        out = torch.empty(n, n, device=DEV)
        fb = _lib.gram_workspace_bytes(n, n, d, True, False)
        wf = torch.empty(fb // 4, device=DEV)
        for lower in (0, 1):
            for _ in range(2):
                call("gram_fwd", 0, ptr(rows), ptr(rows), ptr(sc), d, ptr(var), ptr(out), n, n, n, d, 1e-5, None, lower, ptr(wf), fb)
        if n == 4200:
            bb = _lib.gram_workspace_bytes(n, n, d, True, True, False)
            wb = torch.empty(bb // 4, device=DEV)
            dk = torch.randn(n, n, device=DEV)
            dk = 0.5 * (dk + dk.t())
            dx, dsc, dvar = torch.empty(n, d, device=DEV), torch.empty(d, device=DEV), torch.empty(1, device=DEV)
            for _ in range(2):
                call("gram_bwd", 0, ptr(rows), ptr(rows), ptr(sc), d, ptr(var), ptr(dk), n, n, n, d, ptr(dx), None, ptr(dsc), ptr(dvar), 1, ptr(wb), bb)
        torch.cuda.synchronize()
if which in ("all", "sweep"):
    n = 4200
    a = torch.randn(n, 128, device=DEV)
    k = a @ a.t() / 128 + torch.eye(n, device=DEV)
    ld = ops._pad4(n)
    e0 = torch.zeros(2 * n + 1, ld, device=DEV)
    e0[:n, :n] = k
    e0[n, :n] = torch.randn(n, device=DEV)
    e0[n + 1:, :n].diagonal().fill_(1.0)
    nbytes = _lib.potrf_workspace_bytes(n, 1)
    ws = torch.empty(nbytes // 4, device=DEV)
    info = torch.zeros(1, dtype=torch.int32, device=DEV)
    for _ in range(2):
        e = e0.clone()
        call("potrf_ext", ptr(e), ld, n, 1, 1, None, ptr(ws), nbytes, ptr(info))
    torch.cuda.synchronize()
    assert int(info) == 0
if which in ("all", "trsm"):
    n, r = 4096, 32768
    a = torch.randn(n, 128, device=DEV)
    l = torch.linalg.cholesky(a @ a.t() / 128 + torch.eye(n, device=DEV)).contiguous()
    xx = torch.randn(r, n, device=DEV)
    for _ in range(2):
        call("trsm_rlt", ptr(l), n, n, ptr(xx), n, r)
    torch.cuda.synchronize()
print("ok")
