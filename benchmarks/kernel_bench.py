"""Per-kernel timings at sizes where each kernel meets its own roofline bound (one JSON line each).

    python benchmarks/kernel_bench.py [--quick]

CUDA events on the launching stream, 3 warm-ups, inputs larger than L2 (or L2 flushed) between
launches.  Peaks from MEASURED_PEAKS.json when present."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import pinot_b200 as pb  # noqa: E402
from pinot_b200 import _lib, ops, synthetic  # noqa: E402

DEV = "cuda"


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d["hbm_gbs"], d["bf16_tflops"], "measured"
    return 6650.0, 1590.0, "fallback"


def timeit(fn, iters=10, warm=3, flush=None):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        if flush is not None:
            flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    quick = "--quick" in sys.argv
    only_mp = "--only-mp" in sys.argv
    hbm, tf, src = peaks()
    flush = torch.empty(64 * 2 ** 20, dtype=torch.float32, device=DEV)
    out = []

    # (1) message passing at an HBM-sized batch: 125k molecules = one GPU's share of the 1M pool
    n_mol = 20000 if quick else 125000
    g, _ = synthetic.make_batch(n_mol, seed=1)
    g = g.to(DEV)
    csr = g.csr()
    nn_, ne_ = g.number_of_nodes(), g.number_of_edges()
    for h in (128, 32):
        x = torch.randn(nn_, h, device=DEV)
        for norm in ("both", "mean"):
            ms = timeit(lambda: ops.aggregate(x, csr, norm))
            b = 8.0 * h * nn_ + 4.0 * ne_ + 4.0 * (nn_ + 1) + 8.0 * nn_
            out.append({"kernel": "mp_aggregate_fwd", "H": h, "norm": norm, "molecules": n_mol, "nodes": nn_,
                        "ms": ms, "GBps": b / ms / 1e6, "frac_hbm": b / ms / 1e6 / hbm,
                        "molecules_per_s": n_mol / ms * 1e3})
        ms = timeit(lambda: ops.segment_readout(x, csr))
        b = 4.0 * h * nn_ + 4.0 * h * n_mol
        out.append({"kernel": "segment_readout_fwd", "H": h, "ms": ms, "GBps": b / ms / 1e6,
                    "frac_hbm": b / ms / 1e6 / hbm})
    del x
    if only_mp:
        for o in out:
            print(json.dumps(o))
        return

    # (2) node-update GEMM
    for (m, k, n) in [(nn_ if quick else 1000000, 128, 128), (8192, 8192, 8192) if not quick else (2048, 2048, 2048)]:
        a = torch.randn(m, k, device=DEV)
        w = torch.randn(k, n, device=DEV)
        ms = timeit(lambda: ops.matmul(a, w), flush=flush)
        out.append({"kernel": "sgemm NN", "M": m, "N": n, "K": k, "ms": ms, "TFLOPs": 2.0 * m * n * k / ms / 1e9,
                    "frac_bf16_peak": 2.0 * m * n * k / ms / 1e9 / tf})
        del a, w

    # (2b) node update proper: tanh(X W^T + b), one row per atom, HBM roofline = read X + write Y
    for (m, k, n) in [(nn_, 128, 128)] + ([] if quick else [(1000000, 128, 128), (4000000, 128, 128)]):
        xx = torch.randn(m, k, device=DEV)
        ww = torch.randn(n, k, device=DEV) * 0.1
        bb = torch.randn(n, device=DEV)
        ms = timeit(lambda: ops.node_update(xx, ww, bb, w_is_out_in=True, act="tanh"), flush=flush)
        byts = 4.0 * m * (k + n) + 4.0 * k * n
        out.append({"kernel": "node_update_fwd (tanh)", "rows": m, "in": k, "out": n, "ms": ms, "GBps": byts / ms / 1e6,
                    "frac_hbm": byts / ms / 1e6 / hbm})
        del xx, ww, bb

    # (3) Gram forward / backward, (4) sweep and the full exact-GP fit
    for n in ([1128, 4200] if quick else [1128, 4200, 16384]):
        d = 128
        x = (0.1 * torch.randn(n, d, device=DEV))
        y = torch.randn(n, 1, device=DEV)
        sc, var, ls = torch.zeros(d, device=DEV), torch.zeros((), device=DEV), torch.tensor(-3.0, device=DEV)
        ms = timeit(lambda: ops.gram("rbf", x, None, sc, var), flush=flush)
        out.append({"kernel": "gram_fwd", "N": n, "D": d, "ms": ms, "write_GBps": 4.0 * n * n / ms / 1e6,
                    "frac_hbm": 4.0 * n * n / ms / 1e6 / hbm, "ALU_TFLOPs": 3.0 * n * n * d / ms / 1e9})
        k = ops.gram("rbf", x, None, sc, var, diag_add=1e-5 + 0.05)
        ms = timeit(lambda: ops.cholesky(k.clone() if False else k), iters=5, flush=flush)
        out.append({"kernel": "potrf (sweep, no extension rows)", "N": n, "ms": ms,
                    "TFLOPs": n ** 3 / 3.0 / ms / 1e9, "frac_bf16_peak": n ** 3 / 3.0 / ms / 1e9 / tf})
        # the loss's sweep: stacked [K; y^T; I] (factor + z + L^-T), raw entry point
        from pinot_b200 import _lib
        from pinot_b200._lib import call, ptr
        ld = ops._pad4(n)
        e0 = torch.zeros(2 * n + 1, ld, device=DEV)
        e0[:n, :n] = k
        e0[n, :n] = y.flatten()
        e0[n + 1:, :n].diagonal().fill_(1.0)
        e = e0.clone()
        nbytes = _lib.potrf_workspace_bytes(n, 1)
        ws = torch.empty(nbytes // 4, device=DEV)
        info = torch.zeros(1, dtype=torch.int32, device=DEV)

        def sweep():
            call("potrf_ext", ptr(e), ld, n, 1, 1, None, ptr(ws), nbytes, ptr(info))

        def restore():
            e.copy_(e0)

        ts = []
        for _ in range(6):
            restore()
            if flush is not None:
                flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            sweep()
            b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        ms = sorted(ts[1:])[len(ts[1:]) // 2]
        out.append({"kernel": "potrf_ext [K; y; I] (factor + L^-1 y + L^-T: the loss's sweep)", "N": n, "ms": ms,
                    "TFLOPs": (2.0 * n ** 3 / 3.0 + n * n) / ms / 1e9, "frac_bf16_peak": (2.0 * n ** 3 / 3.0 + n * n) / ms / 1e9 / tf})
        del e, e0, ws
        dk = torch.randn(n, n, device=DEV)
        dk = 0.5 * (dk + dk.t())
        ms = timeit(lambda: ops.gram_bwd_raw("rbf", x, x, sc, var.reshape(1), dk, need_y=False, symmetric=True), iters=5, flush=flush)
        out.append({"kernel": "gram_bwd (symmetric mode, as in the loss)", "N": n, "D": d, "ms": ms,
                    "read_GBps": 4.0 * n * n / ms / 1e6, "frac_hbm": 4.0 * n * n / ms / 1e6 / hbm})
        del dk
        del k
        xg = x.clone().requires_grad_(True)
        lsg = ls.clone().requires_grad_(True)

        def fit():
            xg.grad = None
            lsg.grad = None
            ops.exact_gp_loss(xg, y, sc, var, lsg).sum().backward()

        ms = timeit(fit, iters=5, flush=flush)
        out.append({"kernel": "exact-GP fit (Gram + sweep + alpha + NLL + gradients)", "N": n, "D": d, "ms": ms,
                    "TFLOPs": (n ** 3 + 2.0 * n * n * d) / ms / 1e9})
    for o in out:
        o["peaks"] = src
        print(json.dumps(o))


if __name__ == "__main__":
    main()
