"""How many pairs of the C2 embedding matrix the tensor-core Gram kernel recomputes in difference form, and what
that costs, for the threshold set by PINOT_B200_GRAM_TAU (read once per process: run once per value).
Embeddings: the bench network (GraphConv 3x128, random init) over the bench's 4 200 synthetic molecules, kernel
log length-scale -3 as in the parity tests."""
import json
import os
import sys
import torch
sys.path.insert(0, ".")
os.environ["PINOT_B200_GRAM_DEBUG"] = "1"
import pinot_b200 as pb
from pinot_b200 import _lib, synthetic
from pinot_b200._lib import call, ptr

DEV = "cuda"
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4200
torch.manual_seed(0)
g, y = synthetic.make_batch(n, seed=2666)
rep = pb.representation.Sequential(pb.representation.dgl_legacy.gn("GraphConv"), [128, "tanh", 128, "tanh", 128, "tanh"])
rep = rep.to(DEV)
with torch.no_grad():
    x = rep(g.to(DEV)).contiguous()
d = x.shape[1]
sc = torch.full((d,), -3.0, device=DEV)
var = torch.zeros(1, device=DEV)
out = torch.empty(n, n, device=DEV)
fb = _lib.gram_workspace_bytes(n, n, d, True, False)
wf = torch.zeros(fb // 4, device=DEV)


def fwd():
    call("gram_fwd", 0, ptr(x), ptr(x), ptr(sc), d, ptr(var), ptr(out), n, n, n, d, 1e-5, None, 0, ptr(wf), fb)


evs = []
for _ in range(25):
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record(); fwd(); e.record(); evs.append((s, e))
torch.cuda.synchronize()
ts = sorted(a.elapsed_time(b) for a, b in evs[5:])
bb = _lib.gram_workspace_bytes(n, n, d, True, True, False)
wb = torch.zeros(bb // 4, device=DEV)
dk = torch.randn(n, n, device=DEV)
dk = 0.5 * (dk + dk.t())
dx, dsc, dvar = torch.empty(n, d, device=DEV), torch.empty(d, device=DEV), torch.empty(1, device=DEV)


def bwd():
    call("gram_bwd", 0, ptr(x), ptr(x), ptr(sc), d, ptr(var), ptr(dk), n, n, n, d, ptr(dx), None, ptr(dsc), ptr(dvar), 1, ptr(wb), bb)


evb = []
for _ in range(25):
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record(); bwd(); e.record(); evb.append((s, e))
torch.cuda.synchronize()
tb = sorted(a.elapsed_time(b) for a, b in evb[5:])
xs = (x - x.mean(0)) * sc.exp()
nrm = (xs * xs).sum(1)
dist = torch.cdist(xs.double(), xs.double())
# where the flagged pairs sit: per epilogue-warp piece (32 rows x 64 columns) and per CTA (tile t -> CTA t % 148,
# tiles on or below the diagonal only), from the fp64 distances and the kernel's criterion d < tau (|x|^2 + |y|^2)
tau = float(os.environ.get("PINOT_B200_GRAM_TAU", "0.03"))
flag = (dist < tau * (nrm[:, None] + nrm[None, :]).double()) & (torch.arange(n, device=DEV)[:, None] > torch.arange(n, device=DEV)[None, :])
npad = (n + 127) // 128 * 128
fp = torch.zeros(npad, npad, device=DEV)
fp[:n, :n] = flag.float()
piece = fp.reshape(npad // 32, 32, npad // 64, 64).sum((1, 3))          # (row blocks of 32, column blocks of 64)
tile = fp.reshape(npad // 128, 128, npad // 128, 128).sum((1, 3))
gt = npad // 128
cta = torch.zeros(148, device=DEV)
cta_piece_max = torch.zeros(148, device=DEV)
for i in range(gt):
    for j in range(i + 1):
        t = i * gt + j
        cta[t % 148] += tile[i, j]
        cta_piece_max[t % 148] += piece[4 * i:4 * i + 4, 2 * j:2 * j + 2].max()
print(json.dumps({"n": n, "d": d, "tau": os.environ.get("PINOT_B200_GRAM_TAU", "default"),
                  "fwd_ms_median": ts[len(ts) // 2], "bwd_sym_ms_median": tb[len(tb) // 2],
                  "exact_recomputations_per_call": float(wf[0]) / 25, "pairs_lower_incl_diag_tiles": n * (n + 128) // 2,
                  "median_sq_norm_scaled_centred": float(nrm.median()), "median_distance": float(dist.median()),
                  "flagged_by_fp64_criterion": float(flag.sum()), "piece_max": float(piece.max()),
                  "cta_total_mean": float(cta.mean()), "cta_total_max": float(cta.max()),
                  "cta_sum_of_busiest_piece_per_tile_max": float(cta_piece_max.max()),
                  "frac_pairs_d_below_0p1": float((dist < 0.1).double().mean())}))
