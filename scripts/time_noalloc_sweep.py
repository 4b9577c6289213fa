"""pinot_b200_potrf_ext_noalloc (plain two-launch-per-panel schedule, built-in scratch) against pinot_b200_potrf_ext
(look-ahead, fused factor and panel solve, caller workspace) on [K; y; I].  One JSON line per size."""
import json
import sys
import torch
sys.path.insert(0, ".")
from pinot_b200 import _lib, ops
from pinot_b200._lib import call, ptr

DEV = "cuda"
wa = torch.randn(8192, 8192, device=DEV)
for _ in range(40):
    wa @ wa
for n in [int(a) for a in sys.argv[1:]] or [1128, 4200, 8192]:
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
    diag = torch.empty(n, device=DEV)
    out = {"n": n, "rows": 2 * n + 1}
    for name in ("potrf_ext", "potrf_ext_noalloc"):
        e = e0.clone()
        evs = []
        for _ in range(10):
            e.copy_(e0)
            s, t = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            if name == "potrf_ext":
                call(name, ptr(e), ld, n, 1, 1, ptr(diag), ptr(ws), nbytes, ptr(info))
            else:
                call(name, ptr(e), ld, n, 1, 1, ptr(diag), ptr(info))
            t.record()
            evs.append((s, t))
        torch.cuda.synchronize()
        assert int(info) == 0
        ts = sorted(s.elapsed_time(t) for s, t in evs[3:])
        out[name + "_ms"] = ts[len(ts) // 2]
    out["ratio"] = out["potrf_ext_noalloc_ms"] / out["potrf_ext_ms"]
    print(json.dumps(out))
