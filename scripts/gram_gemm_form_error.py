"""Where the error of the GEMM-form distance comes from (C2 embeddings, tau = 0 so nothing is recomputed).
Relative error of K against fp64, binned by d / (|x|^2 + |y|^2), for
  kernel      : the tensor-core route as built
  emul_now    : its formula evaluated in fp64 (norms of the full fp32 rows, dot = hi.hi + hi.lo + lo.hi): what is
                left is the dropped lo.lo term and the mismatch between the norms and the dot
  emul_norms  : the same with the norms taken from the same three terms (hi^2 + 2 hi lo)
  emul_4term  : dot and norms with the lo.lo term as well
The difference between `kernel` and `emul_now` is accumulation rounding inside the tensor core."""
import json
import os
import sys
import torch
sys.path.insert(0, ".")
os.environ["PINOT_B200_GRAM_TAU"] = "0"
import pinot_b200 as pb
from pinot_b200 import _lib, synthetic
from pinot_b200._lib import call, ptr

DEV = "cuda"
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4200
torch.manual_seed(0)
g, y = synthetic.make_batch(n, seed=2666)
rep = pb.representation.Sequential(pb.representation.dgl_legacy.gn("GraphConv"), [128, "tanh", 128, "tanh", 128, "tanh"]).to(DEV)
with torch.no_grad():
    x = rep(g.to(DEV)).contiguous()
d = x.shape[1]
sc = torch.full((d,), -3.0, device=DEV)
var = torch.zeros(1, device=DEV)
out = torch.empty(n, n, device=DEV)
fb = _lib.gram_workspace_bytes(n, n, d, True, False)
wf = torch.zeros(fb // 4, device=DEV)
call("gram_fwd", 0, ptr(x), ptr(x), ptr(sc), d, ptr(var), ptr(out), n, n, n, d, 0.0, None, 0, ptr(wf), fb)
torch.cuda.synchronize()

xs64 = x.double() * sc.double().exp()
dist = torch.cdist(xs64, xs64)
k64 = torch.exp(-0.5 * dist)

# the kernel's operands: centred by the column mean, scaled, all in fp32; hi = tf32(v) (round to nearest even on
# 13 dropped bits), lo = tf32(v - hi)
v = ((x - x.mean(0)) * sc.exp()).float()


def tf32(t):
    b = t.view(torch.int32)
    b = (b + 0x0FFF + ((b >> 13) & 1)) & ~0x1FFF
    return b.view(torch.float32)


hi = tf32(v)
lo = tf32(v - hi)
H, L, V = hi.double(), lo.double(), v.double()
n_full = (v * v).sum(1).double()                       # fp32 sum, as the prep kernel does
dot3 = H @ H.t() + H @ L.t() + L @ H.t()
n3 = (H * H + 2 * H * L).sum(1)
dot4 = dot3 + L @ L.t()
n4 = ((H + L) ** 2).sum(1)


def k_of(nrm, dot):
    d2 = (nrm[:, None] + nrm[None, :] - 2 * dot).clamp_min(0)
    return torch.exp(-0.5 * d2.sqrt())


cands = {"kernel": out.double(), "emul_now": k_of(n_full, dot3), "emul_norms": k_of(n3, dot3), "emul_4term": k_of(n4, dot4)}
s = n_full[:, None] + n_full[None, :]
ratio = dist / s
off = ~torch.eye(n, dtype=torch.bool, device=DEV)
edges = [0.0, 0.01, 0.03, 0.06, 0.1, 0.2, 0.5, 1e9]
res = {"n": n, "bins_d_over_s": edges, "pairs": [], "max_rel_err": {k: [] for k in cands}, "rms_rel_err": {k: [] for k in cands}}
for a, b in zip(edges[:-1], edges[1:]):
    m = off & (ratio >= a) & (ratio < b)
    res["pairs"].append(int(m.sum()))
    for k, c in cands.items():
        e = ((c - k64).abs() / k64)[m]
        res["max_rel_err"][k].append(float(e.max()) if e.numel() else 0.0)
        res["rms_rel_err"][k].append(float((e * e).mean().sqrt()) if e.numel() else 0.0)
print(json.dumps(res))
