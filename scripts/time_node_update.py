"""Node-update forward at HBM-sized inputs (the node lines of benchmarks/kernel_bench.py plus H = 32 and the plain
variant), same method: CUDA events, 3 warm-ups, L2 flushed between launches, median of 10.  One JSON line each."""
import json, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "benchmarks"))
from kernel_bench import peaks, timeit, DEV   # noqa: E402
from pinot_b200 import ops                    # noqa: E402

hbm, _, src = peaks()
flush = torch.empty(64 * 2 ** 20, dtype=torch.float32, device=DEV)
torch.manual_seed(0)
for (m, k, n) in [(3122861, 128, 128), (1000000, 128, 128), (105316, 128, 128), (3122861, 32, 32), (3122861, 128, 32), (3122861, 32, 128)]:
    xx = torch.randn(m, k, device=DEV)
    ww = torch.randn(n, k, device=DEV) * 0.1
    bb = torch.randn(n, device=DEV)
    byts = 4.0 * m * (k + n) + 4.0 * k * n
    for act in ("tanh", None):
        ms = timeit(lambda: ops.node_update(xx, ww, bb, w_is_out_in=True, act=act), flush=flush)
        print(json.dumps({"kernel": "node_update_fwd (%s)" % (act or "no activation"), "rows": m, "in": k, "out": n, "ms": ms,
                          "GBps": byts / ms / 1e6, "frac_hbm": byts / ms / 1e6 / hbm, "peaks": src}), flush=True)
    del xx, ww, bb
