"""Timing probes for the Cholesky sweep building blocks (CUDA events, back-to-back launches)."""
import os
import sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pinot_b200 import _lib, ops  # noqa: E402
from pinot_b200._lib import call, ptr  # noqa: E402
DEV = "cuda"


def time_ms(fn, reps=5):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record(); torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    return sorted(ts)[len(ts) // 2]


def main():
    print("backend:", os.environ.get("PINOT_B200_GEMM", "tc"), "lookahead:", os.environ.get("PINOT_B200_LOOKAHEAD", "1"))
    big = torch.randn(8500, 4200, device=DEV)
    w = torch.randn(64, 64, device=DEV)
    for (m, n, k, tb) in [(4201, 64, 64, False), (4201, 128, 64, True), (4201, 1024, 64, True), (4201, 2048, 64, True),
                          (4201, 4096, 64, True), (4201, 2048, 128, True), (4201, 2048, 256, True), (105316, 128, 128, False)]:
        if m > 8500:
            a = torch.randn(m, k, device=DEV); b = torch.randn(k, n, device=DEV); c = torch.empty(m, n, device=DEV)
            f = lambda: [call("sgemm", 0, 0, m, n, k, 1.0, ptr(a), k, ptr(b), n, 0.0, ptr(c), n) for _ in range(20)]
        elif not tb:
            f = lambda: [call("sgemm", 0, 0, m, n, k, 1.0, ptr(big), 4200, ptr(w), 64, 0.0, ptr(big), 4200) for _ in range(20)]
        else:
            # C (m x n) -= A (m x k) B(n x k)^T, all views into one big buffer like the sweep
            f = lambda: [call("sgemm", 0, 1, m, n, k, -1.0, ptr(big), 4200, ptr(big), 4200, 1.0, ptr(big[:, 64:]), 4200)
                         for _ in range(20)]
        ms = time_ms(f) / 20
        print("gemm m=%6d n=%5d k=%4d tb=%d : %8.1f us/launch  %7.2f TFLOP/s" % (m, n, k, tb, ms * 1e3, 2.0 * m * n * k / ms / 1e9))
    for n in (1128, 4200, 8192, 16384):
        x = 0.1 * torch.randn(n, 128, device=DEV)
        sc, var = torch.zeros(128, device=DEV), torch.zeros((), device=DEV)
        k = ops.gram("rbf", x, None, sc, var, diag_add=0.05)
        ld = ops._pad4(n)
        for ident in (False, True):
            e = torch.zeros((n + 1 + (n if ident else 0), ld), device=DEV)

            def run():
                e[:n, :n] = k
                if ident:
                    e[n + 1:, :n].zero_()
                    e[n + 1:, :n].diagonal().fill_(1.0)
                ops.potrf_ext_raw(e, n, n_ext_dense=1, ext_identity=ident, check=False)
            def prep():
                e[:n, :n] = k
                if ident:
                    e[n + 1:, :n].zero_()
                    e[n + 1:, :n].diagonal().fill_(1.0)
            t_all, t_prep = time_ms(run), time_ms(prep)
            print("potrf_ext n=%d identity=%d : %.3f ms (prep %.3f ms excluded)" % (n, ident, t_all - t_prep, t_prep))
    # host-side enqueue cost of one tiny launch
    import time
    a = torch.randn(64, 64, device=DEV); c = torch.empty(64, 64, device=DEV)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(2000):
        call("sgemm", 0, 0, 64, 64, 64, 1.0, ptr(a), 64, ptr(a), 64, 0.0, ptr(c), 64)
    t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
    print("host enqueue per sgemm call: %.2f us; drained after %.2f us more per call" % ((t1 - t0) / 2000 * 1e6, (t2 - t1) / 2000 * 1e6))


if __name__ == "__main__":
    main()
