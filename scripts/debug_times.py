import os, sys, ctypes, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pinot_b200 import ops, _lib
DEV = "cuda"
lib = _lib.load()
lib.pinot_b200_debug_times.argtypes = [ctypes.c_void_p]
for n in (330, 1128):
    x = 0.1 * torch.randn(n, 128, device=DEV)
    sc, var = torch.zeros(128, device=DEV), torch.zeros((), device=DEV)
    k = ops.gram("rbf", x, None, sc, var, diag_add=0.05)
    ld = ops._pad4(n)
    e = torch.zeros((2 * n + 1, ld), device=DEV)
    for it in range(3):
        e.zero_(); e[:n, :n] = k; e[n + 1:, :n].diagonal().fill_(1.0)
        ops.potrf_ext_raw(e, n, n_ext_dense=1, ext_identity=True, check=False)
    torch.cuda.synchronize()
    buf = (ctypes.c_ulonglong * 16)()
    lib.pinot_b200_debug_times(buf)
    t = list(buf)[:8]
    names = ["start", "setup done", "mma done", "epilogue done", "postop begin", "factor done", "flag published", "solve done", "solve: loaded", "solve: computed", "solve: staged", "epi c0: tmem loaded", "epi c0: staged", "epi c0: stored", "epi c1: stored"]; t = list(buf)[:15]
    print("n=%d last GEMM with post-ops, tile (0,0), ns relative to start:" % n)
    for nm, v in zip(names, t):
        print("   %-16s %8d" % (nm, v - t[0]))
