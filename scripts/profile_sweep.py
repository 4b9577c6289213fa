import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pinot_b200 import ops
DEV = "cuda"
for n in (1128, 4200):
    x = 0.1 * torch.randn(n, 128, device=DEV)
    sc, var = torch.zeros(128, device=DEV), torch.zeros((), device=DEV)
    k = ops.gram("rbf", x, None, sc, var, diag_add=0.05)
    ld = ops._pad4(n)
    e = torch.zeros((2 * n + 1, ld), device=DEV)
    for it in range(2):
        e.zero_(); e[:n, :n] = k; e[n + 1:, :n].diagonal().fill_(1.0)
        ops.potrf_ext_raw(e, n, n_ext_dense=1, ext_identity=True, check=False)
    torch.cuda.synchronize()
print("done")
