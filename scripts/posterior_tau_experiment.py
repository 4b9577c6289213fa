"""Which Gram matrix the posterior's accuracy depends on: K_tr,tr (factorised) or K_te,tr (solved against the
factor).  C1 shape, both kernels; marginal variance and mean against the fp64 oracle, element-wise (floor 1e-2),
with the near-pair threshold of each of the two Gram calls set separately (PINOT_B200_GRAM_TAU is read per call)."""
import json
import os
import sys
import torch
sys.path.insert(0, ".")
sys.path.insert(0, "tests")
import test_config_parity_gpu as T
from pinot_b200 import ops, synthetic
from pinot_b200.regressors import ExactGaussianProcessRegressor

DEV = "cuda"
for kernel in ("rbf", "matern52"):
    torch.manual_seed(0)
    g, y = synthetic.make_batch(1128, seed=2666)
    g_te = synthetic.make_batch(int(sys.argv[1]) if len(sys.argv) > 1 else 512, seed=4242)[0]   # >= 256: K_te,tr on the tensor cores
    net = T.make_net("GraphConv", ExactGaussianProcessRegressor, kernel=kernel)
    with torch.no_grad():
        net.output_regressor.kernel.scale.fill_(-3.0)
    state = {k: v.detach().clone() for k, v in net.state_dict().items()}
    ok = "matern52_safe" if kernel == "matern52" else "rbf"
    o32 = T.oracle_exact(g, y, g_te, state, ok, torch.float32)
    o64 = T.oracle_exact(g, y, g_te, state, ok, torch.float64)
    net = net.to(DEV)
    with torch.no_grad():
        h = net.representation(g.to(DEV))
        h_te = net.representation(g_te.to(DEV))
    head = net.output_regressor
    kind = "matern52" if kernel == "matern52" else "rbf"

    def err(a, ref):
        a, ref = a.double().cpu(), ref.double()
        return float(((a - ref).abs() / ref.abs().clamp_min(1e-2)).max())

    line = {"kernel": kernel, "fp32_oracle": {"mean": err(o32["mean"], o64["mean"]), "var": err(o32["var"], o64["var"])}}
    for t_tr in ("0.03", "0.1", "ffma"):
        for t_te in ("0.03", "0.06", "0.1", "ffma"):
            def setenv(t):
                if t == "ffma":
                    os.environ["PINOT_B200_GRAM"] = "ffma"
                else:
                    os.environ.pop("PINOT_B200_GRAM", None)
                    os.environ["PINOT_B200_GRAM_TAU"] = t
            setenv(t_tr)
            post = ops.ExactGPPosterior(h, y.to(DEV), head.kernel.scale, head.kernel.variance, head.log_sigma, kind=kind)
            setenv(t_te)
            mean, var = post.mean_var(h_te)
            line["tr=%s te=%s" % (t_tr, t_te)] = {"mean": err(mean, o64["mean"]), "var": err(var, o64["var"])}
    print(json.dumps(line))
