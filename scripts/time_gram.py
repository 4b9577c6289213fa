"""Times the Gram kernels (forward, lower-only forward, symmetric backward) on both routes: tensor cores
(workspace given) and CUDA cores (no workspace).  One JSON line per size."""
import json
import sys
import torch
sys.path.insert(0, ".")
from pinot_b200 import _lib, ops
from pinot_b200._lib import call, ptr

DEV = "cuda"
flush = torch.empty(64 * 2 ** 20, device=DEV)
wa = torch.randn(8192, 8192, device=DEV)
for _ in range(40):
    wa @ wa


def timed(fn, reps=12):
    evs = []
    for _ in range(reps):
        flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        evs.append((s, e))
    torch.cuda.synchronize()
    ts = sorted(s.elapsed_time(e) for s, e in evs[3:])
    return ts[len(ts) // 2]


for n in [int(a) for a in sys.argv[1:]] or [1128, 4200, 16384]:
    d = 128
    g = torch.Generator().manual_seed(n)
    x = (3.0 + torch.randn(n, d, generator=g)).to(DEV)
    sc = torch.full((d,), -3.0, device=DEV)
    var = torch.zeros(1, device=DEV)
    out = torch.empty(n, n, device=DEV)
    dk = torch.randn(n, n, generator=g).to(DEV)
    dk = 0.5 * (dk + dk.t())
    dx = torch.empty(n, d, device=DEV)
    dsc = torch.empty(d, device=DEV)
    dvar = torch.empty(1, device=DEV)
    line = {"n": n, "d": d}
    for route in ("tensor", "cuda_cores"):
        fb = _lib.gram_workspace_bytes(n, n, d, True, False) if route == "tensor" else 0
        bb = _lib.gram_workspace_bytes(n, n, d, True, True, False) if route == "tensor" else 0
        wf = torch.empty(max(fb, 4) // 4, device=DEV)
        wb = torch.empty(max(bb, 4) // 4, device=DEV)

        def fwd(lower):
            call("gram_fwd", 0, ptr(x), ptr(x), ptr(sc), d, ptr(var), ptr(out), n, n, n, d, 1e-5, None, lower,
                 ptr(wf) if fb else None, fb)

        def bwd():
            call("gram_bwd", 0, ptr(x), ptr(x), ptr(sc), d, ptr(var), ptr(dk), n, n, n, d, ptr(dx), None, ptr(dsc), ptr(dvar), 1,
                 ptr(wb) if bb else None, bb)

        line[route] = {"fwd_ms": timed(lambda: fwd(0)), "fwd_lower_ms": timed(lambda: fwd(1)), "bwd_sym_ms": timed(bwd),
                       "workspace_MB": (fb / 2 ** 20, bb / 2 ** 20)}
    t = line["tensor"]
    line["fwd_write_GBps"] = 4.0 * n * n / t["fwd_ms"] / 1e6
    line["fwd_gemm_tflops_fp32_equiv"] = 2.0 * n * n * d / t["fwd_ms"] / 1e9
    print(json.dumps(line), flush=True)
