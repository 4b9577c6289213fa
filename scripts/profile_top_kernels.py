"""Launches the kernels the roofline claims are about, a few times each, for `ncu --set full`:
  * tgemm_tma_kernel (TMA-fed tcgen05) as used by pinot_b200_gp_nll_grad at N = 4200 (bench.py's `roofline` kernel),
  * tgemm_kernel<NT> as one rank-64 trailing update of the Cholesky sweep (m = 4201, n = 4136),
  * mp_aggregate_kernel<4> on a 125k-molecule batch (HBM-sized), H = 128,
  * node_gemm_kernel<tanh> (node-update layer) at 105k rows (C2) and 3.1M rows (HBM-sized)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pinot_b200 import ops, synthetic
from pinot_b200._lib import call, ptr
DEV = "cuda"
n = 4200
w = torch.triu(torch.randn(n, n, device=DEV)) * 0.01
diag = torch.rand(n, device=DEV) + 0.5
alpha = torch.randn(n, device=DEV)
g = torch.ones(1, device=DEV)
dk = torch.empty(n, n, device=DEV)
from pinot_b200 import _lib
ws_bytes = _lib.gp_nll_grad_workspace_bytes(n)
ws = torch.empty(max(ws_bytes // 4, 1), device=DEV)
for _ in range(3):
    call("gp_nll_grad", ptr(w), n, n, ptr(diag), ptr(alpha), ptr(g), 0, ptr(dk), n, ptr(ws), ws_bytes)
big = torch.randn(8500, 4200, device=DEV)
for _ in range(3):
    call("sgemm", 0, 1, 4201, 4136, 64, -1.0, ptr(big), 4200, ptr(big), 4200, 1.0, ptr(big[:, 64:]), 4200)
gr, _ = synthetic.make_batch(125000, seed=1)
gr = gr.to(DEV)
x = torch.randn(gr.number_of_nodes(), 128, device=DEV)
csr = gr.csr()
for _ in range(3):
    ops.aggregate(x, csr, "both")
# node-update layer at the C2 size and at an HBM-sized pool (persistent TMA-fed kernel, gemm_node.cu)
for rows in (105316, 3122861):
    xx = torch.randn(rows, 128, device=DEV)
    ww = torch.randn(128, 128, device=DEV) * 0.1
    bb = torch.randn(128, device=DEV)
    for _ in range(3):
        ops.node_update(xx, ww, bb, w_is_out_in=True, act="tanh")
    del xx
torch.cuda.synchronize()
print("ok")
