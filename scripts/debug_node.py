"""Tall-and-skinny GEMM (gemm_node.cu) against fp64 and against the register-staged kernel, with timings."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pinot_b200._lib import call, ptr
DEV = "cuda"


def timeit(fn, iters=20):
    fn(); torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / iters


def rel(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30))


torch.manual_seed(0)
print("NODE_GEMM =", os.environ.get("PINOT_B200_NODE_GEMM", "on"))
for (m, n, k, tb) in [(4096, 128, 128, 1), (105316, 128, 128, 1), (105316, 128, 128, 0), (5000, 32, 128, 0), (7777, 128, 64, 1),
                      (9000, 20, 36, 1), (1000000, 128, 128, 0)]:
    a = torch.randn(m, k, device=DEV)
    b = torch.randn((n, k) if tb else (k, n), device=DEV)
    c = torch.full((m, n), float("nan"), device=DEV)
    call("sgemm", 0, tb, m, n, k, 1.0, ptr(a), k, ptr(b), b.shape[1], 0.0, ptr(c), n)
    want = a.double() @ (b.double().t() if tb else b.double())
    ms = timeit(lambda: call("sgemm", 0, tb, m, n, k, 1.0, ptr(a), k, ptr(b), b.shape[1], 0.0, ptr(c), n))
    gb = 4.0 * m * (k + n) / 1e9
    print("sgemm m=%7d n=%3d k=%3d tb=%d: rel %.2e nan %d | %.3f ms  %.0f GB/s" % (m, n, k, tb, rel(c, want), int(torch.isnan(c).sum()), ms, gb / ms * 1e3), flush=True)
# node update: bias + tanh, accumulate
m, k, n = 105316, 128, 128
x = torch.randn(m, k, device=DEV) * 0.3
w = torch.randn(n, k, device=DEV) * 0.1
bias = torch.randn(n, device=DEV)
y = torch.empty(m, n, device=DEV)
call("node_update_fwd", ptr(x), ptr(w), ptr(bias), ptr(y), m, k, n, 1, 1, 0)
want = torch.tanh(x.double() @ w.double().t() + bias.double())
print("node_update tanh: rel %.2e" % rel(y, want), "| %.3f ms" % timeit(lambda: call("node_update_fwd", ptr(x), ptr(w), ptr(bias), ptr(y), m, k, n, 1, 1, 0)))
y0 = torch.randn(m, n, device=DEV)
y2 = y0.clone()
call("node_update_fwd", ptr(x), ptr(w.t().contiguous()), ptr(bias), ptr(y2), m, k, n, 0, 0, 1)
print("node_update accumulate (in,out) weight: rel %.2e" % rel(y2, y0.double() + x.double() @ w.double().t() + bias.double()))
# the tanh of the epilogue against torch.tanh (CUDA tanhf) on the same pre-activations: bit for bit
pre = torch.empty(m, n, device=DEV)
call("node_update_fwd", ptr(x), ptr(w), ptr(bias), ptr(pre), m, k, n, 1, 0, 0)
call("node_update_fwd", ptr(x), ptr(w), ptr(bias), ptr(y), m, k, n, 1, 1, 0)
print("tanh epilogue vs torch.tanh of the act=0 output: %d of %d differ (max |diff| %.3e)" % (int((y != torch.tanh(pre)).sum()), y.numel(), float((y - torch.tanh(pre)).abs().max())))
