import sys, time, json, torch
sys.path.insert(0, ".")
import pinot_b200 as pb
from pinot_b200 import active, synthetic
dev = torch.device("cuda", 0)
torch.manual_seed(0)
net = pb.Net(pb.representation.Sequential(pb.representation.dgl_legacy.gn(), [128, "tanh", 128, "tanh", 128, "tanh"]),
             output_regressor_class=pb.regressors.ExactGaussianProcessRegressor).to(dev)
with torch.no_grad():
    net.output_regressor.kernel.scale.fill_(-3.0)
g_tr, y_tr = synthetic.make_batch(4096, seed=11, device=dev)
pool = synthetic.make_batch(125000, seed=100)[0].to(dev)
pool.csr()
net.g_last, net.y_last = g_tr, y_tr
def t(fn, reps=3):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        out = fn()
    torch.cuda.synchronize()
    return 1e3 * (time.perf_counter() - t0) / reps, out
ms_round, _ = t(lambda: active.thompson_sampling_sharded(net, pool, 0, q=96))
with torch.no_grad():
    ms_post, post = t(lambda: net._posterior())
    ms_sel, sub = t(lambda: pool.select(range(0, 65536)))
    ms_rep, h = t(lambda: net.representation(sub))
    ms_mv, _ = t(lambda: post.mean_var(h))
    ms_gram, gm = t(lambda: pb.ops.gram_raw(post.kind, h[:32768], post.x_tr, post.scale, post.variance))
    ms_cm, d = t(lambda: net.condition_marginal(pool))
    ms_draw, _ = t(lambda: (d.loc + d.scale * torch.randn((96, d.loc.numel()), device=dev)).max(dim=1))
print(json.dumps({"round_ms": ms_round, "posterior_fetch_ms": ms_post, "select_65536_ms": ms_sel, "embed_65536_ms": ms_rep,
                  "mean_var_65536_ms": ms_mv, "gram_32768x4096_ms": ms_gram, "condition_marginal_125k_ms": ms_cm, "draws_ms": ms_draw}))
