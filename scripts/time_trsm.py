"""trsm_rlt (X <- X L^-T, n = 4096) per number of rows and outer block width (PINOT_B200_TRSM_OB)."""
import os, sys, json, torch
sys.path.insert(0, ".")
from pinot_b200._lib import call, ptr
DEV = "cuda"
n = 4096
a = torch.randn(n, 128, device=DEV)
l = torch.linalg.cholesky(a @ a.t() / 128 + torch.eye(n, device=DEV)).contiguous()
for r in [int(v) for v in sys.argv[1:]] or (4096, 32768):
    x0 = torch.randn(r, n, device=DEV)
    for ob in ("128", "512"):
        os.environ["PINOT_B200_TRSM_OB"] = ob
        x = x0.clone()
        evs = []
        for _ in range(5):
            x.copy_(x0)
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record(); call("trsm_rlt", ptr(l), n, n, ptr(x), n, r); e.record(); evs.append((s, e))
        torch.cuda.synchronize()
        ts = sorted(s.elapsed_time(e) for s, e in evs[1:])
        ms = ts[len(ts) // 2]
        print(json.dumps({"r": r, "n": n, "outer_block": int(ob), "ms": ms, "tflops_fp32_equiv": r * n * n / ms / 1e9}))
    # the entry point with caller scratch: 512-column blocks, wide updates accumulated accurately on the TMA-fed GEMM
    os.environ["PINOT_B200_TRSM_OB"] = "128"
    from pinot_b200 import _lib
    nbytes = _lib.trsm_rlt_workspace_bytes(n, r)
    if nbytes:
        ws = torch.empty(nbytes // 4, device=DEV)
        x = x0.clone()
        evs = []
        for _ in range(5):
            x.copy_(x0)
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record(); call("trsm_rlt_ws", ptr(l), n, n, ptr(x), n, r, ptr(ws), nbytes); e.record(); evs.append((s, e))
        torch.cuda.synchronize()
        ts = sorted(s.elapsed_time(e) for s, e in evs[1:])
        ms = ts[len(ts) // 2]
        print(json.dumps({"r": r, "n": n, "entry": "trsm_rlt_ws", "ms": ms, "tflops_fp32_equiv": r * n * n / ms / 1e9}))
