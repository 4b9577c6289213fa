"""potrf_ext timing at a few sizes (identity extension rows, as in the train step) for the current
PINOT_B200_OUTER_BLOCK setting."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pinot_b200 import ops
DEV = "cuda"


def t(fn, reps=5):
    fn(); torch.cuda.synchronize(); ts = []
    for _ in range(reps):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record(); torch.cuda.synchronize(); ts.append(s.elapsed_time(e))
    return sorted(ts)[len(ts) // 2]


for n in [int(v) for v in (sys.argv[1:] or ["2048", "3072", "4200", "5120", "8192"])]:
    x = 0.1 * torch.randn(n, 128, device=DEV)
    k = ops.gram("rbf", x, None, torch.zeros(128, device=DEV), torch.zeros((), device=DEV), diag_add=0.05)
    ld = ops._pad4(n)
    e = torch.zeros((2 * n + 1, ld), device=DEV)

    def prep():
        e[:n, :n] = k
        e[n + 1:, :n].zero_()
        e[n + 1:, :n].diagonal().fill_(1.0)

    def run():
        prep()
        ops.potrf_ext_raw(e, n, n_ext_dense=1, ext_identity=True, check=False)

    print("outer=%s n=%d: %.3f ms" % (os.environ.get("PINOT_B200_OUTER_BLOCK", "default"), n, t(run) - t(prep)), flush=True)
