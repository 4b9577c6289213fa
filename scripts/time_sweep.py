"""Times pinot_b200_potrf_ext on [K; y; I] with and without the workspace tail (single-launch vs launch-per-panel
below n = 6144; TMA-fed vs register-staged rank-512 updates above).  One JSON line per size."""
import json
import sys
import torch
sys.path.insert(0, ".")
from pinot_b200 import _lib, ops
from pinot_b200._lib import call, ptr

DEV = "cuda"
flush = torch.empty(64 * 2 ** 20, device=DEV)
wa = torch.randn(8192, 8192, device=DEV)
for _ in range(60):      # warm the clocks up: a cold, mostly idle GPU runs these short kernels at a fraction of its clock
    wa @ wa
for n in [int(a) for a in sys.argv[1:]] or [1128, 2048, 4200, 6000]:
    g = torch.Generator().manual_seed(n)
    a = torch.randn(n, 128, generator=g).to(DEV)
    k = a @ a.t() / 128 + 1.0 * torch.eye(n, device=DEV)
    ld = ops._pad4(n)
    for ident in (1, 0):
        rows = n + 1 + (n if ident else 0)
        e0 = torch.zeros(rows, ld, device=DEV)
        e0[:n, :n] = k
        e0[n, :n] = torch.randn(n, generator=g).to(DEV)
        if ident:
            e0[n + 1:, :n].diagonal().fill_(1.0)
        out = {"n": n, "identity_rows": ident}
        import os
        for tail in (False, True):
            os.environ["PINOT_B200_SWEEP"] = "persistent" if tail else "launch"
            nbytes = _lib.potrf_workspace_bytes(n, 1)
            ws = torch.empty(nbytes // 4, device=DEV)
            info = torch.zeros(1, dtype=torch.int32, device=DEV)
            e = e0.clone()
            ts, evs = [], []
            for it in range(24):                  # back to back, no host synchronisation in between
                e.copy_(e0)
                flush.zero_()
                wa[:2048] @ wa[:, :2048]
                s, t = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                s.record()
                call("potrf_ext", ptr(e), ld, n, 1, ident, None, ptr(ws), nbytes, ptr(info))
                t.record()
                evs.append((s, t))
            torch.cuda.synchronize()
            ts = [s.elapsed_time(t) for s, t in evs[4:]]
            ts.sort()
            out["with_tail_ms" if tail else "base_ms"] = ts[len(ts) // 2]
        flop = n ** 3 / 3.0 * (2 if ident else 1) + n * n
        out["with_tail_tflops"] = flop / out["with_tail_ms"] / 1e9
        print(json.dumps(out), flush=True)
