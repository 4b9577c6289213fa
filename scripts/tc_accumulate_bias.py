"""Is the tensor core's fp32 accumulation biased (truncation) and by how much?  3xTF32 product of the C2 embedding
matrix with itself (K = 128: 16 accumulating MMAs of K = 8 per term) against the same three terms in fp64."""
import json
import sys
import torch
sys.path.insert(0, ".")
import pinot_b200 as pb
from pinot_b200 import synthetic
from pinot_b200._lib import call, ptr

DEV = "cuda"
n = 4200
torch.manual_seed(0)
g, y = synthetic.make_batch(n, seed=2666)
rep = pb.representation.Sequential(pb.representation.dgl_legacy.gn("GraphConv"), [128, "tanh", 128, "tanh", 128, "tanh"]).to(DEV)
with torch.no_grad():
    x = rep(g.to(DEV)).contiguous()
d = x.shape[1]
v = ((x - x.mean(0)) * 0.049787068).contiguous()
hi, lo = torch.empty(n, d, device=DEV), torch.empty(n, d, device=DEV)
cnt = torch.zeros(64, dtype=torch.int32, device=DEV)
out = torch.empty(n, n, device=DEV)
call("tf32_split", ptr(v), d, n, d, None, 0, ptr(hi), ptr(lo))
call("sgemm_nt_presplit", n, n, d, 1.0, ptr(hi), ptr(lo), ptr(hi), ptr(lo), 0.0, ptr(out), n, 0, ptr(cnt))
torch.cuda.synchronize()
H, L = hi.double(), lo.double()
ref = H @ H.t() + H @ L.t() + L @ H.t()
err = (out.double() - ref)
big = ref.abs() > 0.05
rel = (err / ref.abs())[big]                 # signed, relative to |a|: truncation toward zero gives sign(a) * rel < 0
sgn = (err * ref.sign() / ref.abs())[big]
print(json.dumps({"pairs": int(big.sum()), "mean_signed_rel_err_toward_zero": float(sgn.mean()), "std_rel_err": float(sgn.std()),
                  "quantiles_signed": [float(q) for q in torch.quantile(sgn[:4000000], torch.tensor([0.01, 0.25, 0.5, 0.75, 0.99], device=DEV, dtype=torch.float64))],
                  "ulp_fp32_rel": 2.0 ** -24}))
