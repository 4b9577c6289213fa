"""Per-entry-point breakdown of one exact-GP fit (loss + backward) at N embeddings (default 16384, C4)."""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pinot_b200 import _lib, ops  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=16384)
    ap.add_argument("--d", type=int, default=128)
    ap.add_argument("--reps", type=int, default=3)
    a = ap.parse_args()
    dev = "cuda"
    gen = torch.Generator(device="cpu").manual_seed(0)
    x = (0.1 * torch.randn(a.n, a.d, generator=gen)).to(dev).requires_grad_(True)
    y = torch.randn(a.n, 1, generator=gen).to(dev)
    sc = torch.zeros(a.d, device=dev, requires_grad=True)
    var = torch.zeros((), device=dev, requires_grad=True)
    ls = torch.tensor(-3.0, device=dev, requires_grad=True)

    def fit():
        for t in (x, sc, var, ls):
            t.grad = None
        ops.exact_gp_loss(x, y, sc, var, ls).sum().backward()

    fit()
    torch.cuda.synchronize()
    names = ["gram_fwd", "potrf_ext", "gemv_rows", "gp_nll", "gp_nll_grad", "gram_bwd"]
    _lib.probe = {k: [] for k in names}
    tot = []
    for _ in range(a.reps):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fit()
        e.record()
        torch.cuda.synchronize()
        tot.append(s.elapsed_time(e))
    out = {"n": a.n, "d": a.d, "fit_ms": sorted(tot)[len(tot) // 2]}
    for k in names:
        ts = [s.elapsed_time(e) for s, e in _lib.probe[k]]
        out[k + "_ms"] = sum(ts) / max(a.reps, 1)
    _lib.probe = None
    out["other_ms"] = out["fit_ms"] - sum(v for k, v in out.items() if k.endswith("_ms") and k != "fit_ms")
    print(json.dumps(out))


if __name__ == "__main__":
    main()
