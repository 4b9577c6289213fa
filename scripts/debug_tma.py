"""Diagnostics for the TMA-fed tcgen05 GEMM (gemm_tma.cu): exact integer cases, random accuracy against
fp64, comparison + timing against the register-staged kernel (pinot_b200_sgemm)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pinot_b200 import _lib  # noqa: E402
from pinot_b200._lib import call, ptr  # noqa: E402

DEV = "cuda"


def ws_for(m, n, k):
    nbytes = int(_lib.load().pinot_b200_sgemm_nt_tma_workspace_bytes(m, n, k))
    return torch.empty(nbytes // 4 + 64, device=DEV), nbytes


def tma(a, b, alpha=1.0, beta=0.0, c=None, ws=None):
    m, k = a.shape
    n = b.shape[0]
    if c is None:
        c = torch.full((m, n), float("nan"), device=DEV)
    if ws is None:
        ws = ws_for(m, n, k)
    call("sgemm_nt_tma", m, n, k, alpha, ptr(a), a.stride(0), ptr(b), b.stride(0), beta, ptr(c), c.stride(0), ptr(ws[0]), ws[1])
    return c


def old(a, b, alpha=1.0, beta=0.0, c=None):
    m, k = a.shape
    n = b.shape[0]
    if c is None:
        c = torch.full((m, n), float("nan"), device=DEV)
    call("sgemm", 0, 1, m, n, k, alpha, ptr(a), a.stride(0), ptr(b), b.stride(0), beta, ptr(c), c.stride(0))
    return c


def report(name, got, want):
    torch.cuda.synchronize()
    got, want = got.double().cpu(), want.double().cpu()
    err = (got - want).abs()
    rel = float((got - want).norm() / want.norm().clamp_min(1e-30))
    nbad = int((err > 1e-3 * want.abs().max()).sum())
    print("%-46s rel %.3e  max abs %.3e  bad %d/%d  nan %d" % (name, rel, float(err.max()), nbad, err.numel(),
                                                            int(torch.isnan(got).sum())), flush=True)
    if nbad and err.numel() <= 256 * 256:
        bad = err > 1e-3 * want.abs().max()
        print("   bad rows (first 20):", bad.any(1).nonzero().flatten().tolist()[:20],
              " bad cols (first 20):", bad.any(0).nonzero().flatten().tolist()[:20])
        print("   got[0,:8]", got[0, :8].tolist())
        print("   want[0,:8]", want[0, :8].tolist())


def timeit(fn, iters=10):
    fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / iters


def main():
    torch.manual_seed(0)
    for (m, n, k) in [(128, 128, 8), (128, 128, 32), (128, 128, 64), (128, 128, 128), (256, 256, 96), (128, 128, 1024),
                      (128, 128, 2048)]:
        a = ((torch.arange(m * k).reshape(m, k) * 7 + 3) % 11 - 5).float().to(DEV)
        b = ((torch.arange(n * k).reshape(n, k) * 5 + 1) % 13 - 6).float().to(DEV)
        report("int m%d n%d k%d" % (m, n, k), tma(a, b), a.double() @ b.double().t())
    a = torch.eye(128, device=DEV)
    b = torch.randn(128, 128, device=DEV)
    report("A=I", tma(a, b), b.t())
    report("B=I", tma(b, a), b)
    for (m, n, k) in [(128, 128, 128), (512, 384, 1000), (4200, 4200, 4200), (1000, 200, 116), (130, 72, 36), (2048, 2048, 2048)]:
        a = torch.randn(m, k, device=DEV)
        b = torch.randn(n, k, device=DEV)
        want = a.double() @ b.double().t()
        report("rand m%d n%d k%d (tma)" % (m, n, k), tma(a, b), want)
        report("rand m%d n%d k%d (reg)" % (m, n, k), old(a, b), want)
        c0 = torch.randn(m, n, device=DEV)
        report("   alpha/beta (tma)", tma(a, b, 0.5, -2.0, c0.clone()), 0.5 * want - 2.0 * c0.double())
    for nn in (2048, 4096, 8192):
        a = torch.randn(nn, nn, device=DEV)
        b = torch.randn(nn, nn, device=DEV)
        c = torch.empty(nn, nn, device=DEV)
        ws = ws_for(nn, nn, nn)
        t_new = timeit(lambda: tma(a, b, c=c, ws=ws))
        t_old = timeit(lambda: old(a, b, c=c))
        fl = 2.0 * nn ** 3 / 1e9
        print("n=%d  tma %.3f ms (%.1f TFLOP/s fp32-equiv)   reg %.3f ms (%.1f)" % (nn, t_new, fl / t_new, t_old, fl / t_old),
              flush=True)


if __name__ == "__main__":
    main()
