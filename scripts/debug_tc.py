"""Diagnostics for the tcgen05 GEMM: exact small-integer cases first (layout / descriptor errors show
up as structured mismatches), then random data (3xTF32 accuracy), then odd shapes and all layouts."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pinot_b200 import _lib  # noqa: E402
from pinot_b200._lib import call, ptr  # noqa: E402

DEV = "cuda"


def sgemm(a, b, ta, tb, m, n, k, alpha=1.0, beta=0.0, c=None):
    if c is None:
        c = torch.full((m, n), float("nan"), device=DEV)
    call("sgemm", int(ta), int(tb), m, n, k, alpha, ptr(a), a.shape[1], ptr(b), b.shape[1], beta, ptr(c), c.shape[1])
    torch.cuda.synchronize()
    return c


def report(name, got, want):
    got, want = got.double().cpu(), want.double().cpu()
    err = (got - want).abs()
    rel = float((got - want).norm() / want.norm().clamp_min(1e-30))
    nbad = int((err > 1e-3 * want.abs().max()).sum())
    print("%-46s rel %.3e  max abs %.3e  bad %d/%d  nan %d" % (name, rel, float(err.max()), nbad, err.numel(),
                                                            int(torch.isnan(got).sum())), flush=True)
    if nbad and err.numel() <= 128 * 128:
        bad = (err > 1e-3 * want.abs().max())
        rows = bad.any(1).nonzero().flatten().tolist()
        cols = bad.any(0).nonzero().flatten().tolist()
        print("   bad rows (first 20):", rows[:20], " bad cols (first 20):", cols[:20])
        print("   got[0,:8]", got[0, :8].tolist())
        print("   want[0,:8]", want[0, :8].tolist())
    return rel


def main():
    print("GEMM backend:", os.environ.get("PINOT_B200_GEMM", "tc (default)"))
    torch.manual_seed(0)
    # exact integer cases, NT (A [m][k], B [n][k])
    for (m, n, k) in [(128, 128, 8), (128, 128, 16), (128, 128, 48), (128, 64, 16), (64, 128, 32), (256, 256, 64)]:
        a = ((torch.arange(m * k).reshape(m, k) * 7 + 3) % 11 - 5).float().to(DEV)
        b = ((torch.arange(n * k).reshape(n, k) * 5 + 1) % 13 - 6).float().to(DEV)
        report("int NT m%d n%d k%d" % (m, n, k), sgemm(a, b, False, True, m, n, k), a.double() @ b.double().t())
    # identity probes: A = I (k = m): result must equal B^T rows
    m = n = k = 128
    a = torch.eye(128, device=DEV)
    b = torch.randn(n, k, device=DEV)
    report("A=I NT", sgemm(a, b, False, True, m, n, k), b.t())
    a = torch.randn(m, k, device=DEV)
    report("B=I NT", sgemm(a, torch.eye(128, device=DEV), False, True, m, n, k), a)
    # random: accuracy of the 3xTF32 split
    for (m, n, k) in [(128, 128, 128), (512, 384, 1000), (4201, 64, 64), (1000, 128, 117)]:
        a = torch.randn(m, k, device=DEV)
        b = torch.randn(n, k, device=DEV)
        report("rand NT m%d n%d k%d" % (m, n, k), sgemm(a, b, False, True, m, n, k), a.double() @ b.double().t())
    for ta in (False, True):
        for tb in (False, True):
            m, n, k = 257, 130, 77
            a = torch.randn((k, m) if ta else (m, k), device=DEV)
            b = torch.randn((n, k) if tb else (k, n), device=DEV)
            want = (a.double().t() if ta else a.double()) @ (b.double().t() if tb else b.double())
            report("rand ta=%d tb=%d m%d n%d k%d" % (ta, tb, m, n, k), sgemm(a, b, ta, tb, m, n, k), want)
    # wide-tile path (M, N >= 1024, K >= 256): exact integers, odd shapes, all layouts
    m, n, k = 1280, 1536, 256
    a = ((torch.arange(m * k).reshape(m, k) * 7 + 3) % 11 - 5).float().to(DEV)
    b = ((torch.arange(n * k).reshape(n, k) * 5 + 1) % 13 - 6).float().to(DEV)
    report("wide int NT m%d n%d k%d" % (m, n, k), sgemm(a, b, False, True, m, n, k), a.double() @ b.double().t())
    for ta in (False, True):
        for tb in (False, True):
            m, n, k = 1100, 1300, 300
            a = torch.randn((k, m) if ta else (m, k), device=DEV)
            b = torch.randn((n, k) if tb else (k, n), device=DEV)
            want = (a.double().t() if ta else a.double()) @ (b.double().t() if tb else b.double())
            report("wide rand ta=%d tb=%d m%d n%d k%d" % (ta, tb, m, n, k), sgemm(a, b, ta, tb, m, n, k), want)
    a = torch.randn(2048, 1000, device=DEV); b = torch.randn(1000, 2048, device=DEV); c0 = torch.randn(2048, 2048, device=DEV)
    report("wide alpha/beta NN k=1000", sgemm(a, b, False, False, 2048, 2048, 1000, alpha=-1.0, beta=1.0, c=c0.clone()),
           c0.double() - a.double() @ b.double())
    # beta / accumulate
    a = torch.randn(300, 200, device=DEV)
    b = torch.randn(200, 150, device=DEV)
    c0 = torch.randn(300, 150, device=DEV)
    report("alpha/beta NN", sgemm(a, b, False, False, 300, 150, 200, alpha=-1.0, beta=1.0, c=c0.clone()),
           c0.double() - a.double() @ b.double())
    # split-K path (weight gradient shape)
    a = torch.randn(20000, 128, device=DEV)
    b = torch.randn(20000, 128, device=DEV)
    report("split-K TN k=20000", sgemm(a, b, True, False, 128, 128, 20000), a.double().t() @ b.double())
    print("launches:", _lib.launch_count())


if __name__ == "__main__":
    main()
