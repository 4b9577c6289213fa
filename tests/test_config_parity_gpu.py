"""Oracle parity AT THE SIZES BASELINE.json names (C1, C2, C3, and the sharded
blocked sweep at its own test size), through the public API and the C-ABI kernels.

The oracle is run twice on the same inputs and weights: in fp32 (the reference's own arithmetic:
oracle/net_ref.py follows pinot/net.py:170-248, pinot/representation/sequential.py:78-97,
pinot/regressors/gaussian_process_regressor.py:85-233 and :346-508) and in fp64 (the exact values of
the same expressions).  A quantity passes when it is within the stated tolerance of the fp32
reference; where the reference's own fp32 rounding noise is larger than that tolerance (long
cancelling sums: hyper-parameter gradients, variances of near-duplicates, the ill-conditioned
variational covariance at n = 4096, the posterior at C1 whose K + sigma^2 I has condition number ~2e4), it
passes when it is as close to the fp64 value as the fp32 reference is, give or take a factor of two (two
fp32 evaluations with different operation orders have independent errors of the same size: "never worse
than the other one" would be a coin flip).  Every comparison prints all three distances, so the log shows
which clause held.

Tolerances (north_star): embeddings, predictive mean and variance 1e-4 relative, NLL 1e-3, gradients
5e-3 norm-wise.  "Relative" is asserted twice: norm-wise at 1e-5 (embeddings) / 1e-4, and element-wise at
1e-4 with every entry measured against max(|reference entry|, FLOOR * largest |reference entry|),
FLOOR = 1e-2.  Why a floor at all: an embedding is a sum of ~25 tanh outputs of magnitude <= 1 per
column, so an entry a hundred times smaller than the largest one is the remainder of a cancellation and
fp32 itself cannot hold it to 1e-4 -- the fp32 reference sits 6e-5 ... 2.4e-4 from the fp64 values under
a 1e-3 floor (measured on the B200 box, round 2), i.e. the reference fails its own bar there.  Each
comparison prints the fp32 reference's own distance from fp64 under the same measure.

All comparisons of a test are evaluated and printed before the test fails, so one GPU run shows the
whole table.
"""
import math

import pytest
import torch

import pinot_b200 as pb
from pinot_b200 import sharded_gp, synthetic
from pinot_b200.regressors import ExactGaussianProcessRegressor, VariationalGaussianProcessRegressor
from pinot_b200.regressors.kernels import Matern52
from conftest import max_rel_err, rel_err
from oracle import gp_ref, net_ref

pytestmark = pytest.mark.gpu
DEV = "cuda"
CONFIG = [128, "tanh", 128, "tanh", 128, "tanh"]
FLOOR = 1e-2   # element-wise relative errors are measured against max(|ref|, FLOOR * max|ref|)
FAILS = []     # comparisons that missed their bar in the running test (asserted empty at its end)


def ref_graph(g):
    s, d = g.edges()
    return net_ref.graph_from_arrays(s, d, g.number_of_nodes(), g.batch_num_nodes().tolist(),
                                     g.batch_num_edges().tolist(), g.ndata["h"])


def close(name, ours, ref32, ref64, tol, elementwise=False):
    """Parity clause described in the module docstring; returns the three distances for the log."""
    err = (lambda a, b: max_rel_err(a, b, floor=FLOOR)) if elementwise else rel_err
    e32 = err(ours, ref32)
    e64 = err(ours, ref64)
    noise = err(ref32, ref64)
    print("%-34s vs fp32 oracle %.2e | vs fp64 oracle %.2e | fp32 oracle vs fp64 %.2e | tol %.0e"
          % (name, e32, e64, noise, tol))
    if not (e32 < tol or e64 <= max(tol, 2.0 * noise)):
        FAILS.append((name, e32, e64, noise, tol))
    return e32, e64, noise


def scalar_close(name, ours, ref32, ref64, tol):
    ours, ref32, ref64 = float(ours), float(ref32), float(ref64)
    scale = max(1.0, abs(ref64))
    e32, e64, noise = abs(ours - ref32) / scale, abs(ours - ref64) / scale, abs(ref32 - ref64) / scale
    print("%-34s %.6f | fp32 oracle %.6f | fp64 oracle %.6f | tol %.0e" % (name, ours, ref32, ref64, tol))
    if not (e32 < tol or e64 <= max(tol, 2.0 * noise)):
        FAILS.append((name, ours, ref32, ref64, tol))


@pytest.fixture(autouse=True)
def _collect_failures():
    FAILS.clear()
    yield
    assert not FAILS, "comparisons outside their bar: %r" % (FAILS,)


def make_net(layer, head, kernel=None, **head_kw):
    kw = dict(head_kw)
    if kernel == "matern52":
        kw["kernel"] = lambda in_features: Matern52()
    rep = pb.representation.Sequential(pb.representation.dgl_legacy.gn(layer), CONFIG)
    return pb.Net(rep, output_regressor_class=head, **kw)


def oracle_exact(g, y, g_te, state, kernel, dtype, chunked=False):
    """Loss, embeddings, all gradients and the predictive mean / covariance from the oracle.
    ``chunked``: the GP half through oracle.gp_ref.exact_gp_fit_chunked (the reference arithmetic with the
    Gram matrix built and differentiated 256 rows at a time: at N = 4 200 the unchunked (N, N, 128) tensors
    need ~80 GB of host memory, which the GPU box does not reliably have free)."""
    p = {k: v.detach().clone().to(dtype).requires_grad_(True) for k, v in state.items()}
    rg = ref_graph(g)
    if chunked:
        assert kernel == "rbf"
        h = net_ref.representation(rg, p, CONFIG)
        gp = "output_regressor."
        loss, gr = gp_ref.exact_gp_fit_chunked(h.detach(), y.to(dtype), p[gp + "kernel.scale"].detach(),
                                               p[gp + "kernel.variance"].detach(), p[gp + "log_sigma"].detach())
        h.backward(gr["x"])
        p[gp + "kernel.scale"].grad, p[gp + "kernel.variance"].grad = gr["scale"], gr["variance"]
        p[gp + "log_sigma"].grad = gr["log_sigma"]
    else:
        loss, h = net_ref.exact_net_loss(rg, y.to(dtype), p, CONFIG, kernel=kernel)
        loss.sum().backward()
    out = {"loss": loss.detach(), "h": h.detach(), "grads": {k: v.grad for k, v in p.items()}}
    if g_te is not None:
        with torch.no_grad():
            q = {k: v.detach() for k, v in p.items()}
            mean, cov = net_ref.exact_net_condition(ref_graph(g_te), rg, y.to(dtype), q, CONFIG, kernel=kernel)
        out["mean"], out["var"] = mean, torch.diagonal(cov).clone()
    return out


def run_exact_config(n_mol, kernel, n_test, fp64=True, chunked=False):
    torch.manual_seed(0)
    g, y = synthetic.make_batch(n_mol, seed=2666)
    g_te = synthetic.make_batch(n_test, seed=4242)[0] if n_test else None
    net = make_net("GraphConv", ExactGaussianProcessRegressor, kernel=kernel)
    with torch.no_grad():
        # "trained-like" length-scale; Matern52 has one scalar scale
        net.output_regressor.kernel.scale.fill_(-3.0)
    state = {k: v.detach().clone() for k, v in net.state_dict().items()}
    okernel = "matern52_safe" if kernel == "matern52" else "rbf"   # identical forward, finite gradients
    o32 = oracle_exact(g, y, g_te, state, okernel, torch.float32, chunked)
    o64 = oracle_exact(g, y, g_te, state, okernel, torch.float64, chunked) if fp64 else o32

    net = net.to(DEV)
    gd, yd = g.to(DEV), y.to(DEV)
    h = net.representation(gd)
    close("embeddings h (element-wise)", h, o32["h"], o64["h"], 1e-4, elementwise=True)
    close("embeddings h (norm-wise)", h, o32["h"], o64["h"], 1e-5)
    net.zero_grad()
    loss = net.loss(gd, yd)
    assert loss.shape == (1, 1) and loss.dtype == torch.float32
    scalar_close("NLL", loss.sum(), o32["loss"].sum(), o64["loss"].sum(), 1e-3)
    loss.sum().backward()
    for k, p in net.named_parameters():
        g32, g64 = o32["grads"][k], o64["grads"][k]
        assert p.grad is not None and torch.isfinite(p.grad).all(), k
        if float(g64.abs().max()) < 1e-7:
            assert float(p.grad.abs().max()) < 1e-5, k
            continue
        close("grad " + k.replace("representation.", "rep.").replace("output_regressor.", "gp."), p.grad, g32, g64, 5e-3)
    if g_te is not None:
        with torch.no_grad():
            dist = net.condition(g_te.to(DEV))
            marg = net.condition_marginal(g_te.to(DEV))
        close("predictive mean (norm-wise)", dist.mean, o32["mean"], o64["mean"], 1e-4)
        close("predictive variance (norm-wise)", dist.variance, o32["var"], o64["var"], 1e-4)
        close("predictive mean (element-wise)", dist.mean, o32["mean"], o64["mean"], 1e-4, elementwise=True)
        close("predictive variance (element-wise)", dist.variance, o32["var"], o64["var"], 1e-4, elementwise=True)
        close("marginal mean (element-wise)", marg.mean, o32["mean"], o64["mean"], 1e-4, elementwise=True)
        close("marginal variance (element-wise)", marg.variance, o32["var"], o64["var"], 1e-4, elementwise=True)


@pytest.mark.parametrize("kernel", ["rbf", "matern52"])
def test_c1_esol_shape_matches_oracle(kernel):
    """configs[0] / C1: 1 128 molecules (25 atoms / 54 directed edges), GraphConv 3x128 + exact GP."""
    run_exact_config(1128, kernel, n_test=200)


def test_c2_4200_molecules_one_step_matches_oracle():
    """configs[1] / C2 (the bench workload): 4 200 molecules, one loss + backward against the oracle in
    fp32 and fp64, the GP half row-chunked (same arithmetic per element, see oracle_exact)."""
    run_exact_config(4200, "rbf", n_test=0, chunked=True)


@pytest.mark.parametrize("n_batch", [64, 4096])
def test_c3_sage_vgp_512_inducing_points_matches_oracle(n_batch):
    """configs[2] / C3: GraphSAGE 3x128 + VariationalGaussianProcessRegressor with M = 512 inducing
    points, minibatches of 64 and 4 096 molecules (pinot/regressors/gaussian_process_regressor.py:346-508)."""
    torch.manual_seed(1)
    g, y = synthetic.make_batch(n_batch, seed=77)
    g_te, _ = synthetic.make_batch(96, seed=78)
    net = make_net("SAGEConv", VariationalGaussianProcessRegressor, n_inducing_points=512)
    with torch.no_grad():
        net.output_regressor.kernel.scale.fill_(-3.0)
        net.output_regressor.y_tr_sigma_tril.normal_(0.0, 0.02)   # non-trivial strictly-lower scale
    state = {k: v.detach().clone() for k, v in net.state_dict().items()}
    # Inducing inputs inside the cloud of embeddings: the initial U(-1, 1) box sits far from the
    # sum-pooled embeddings, which would make K_uf numerically zero and the test vacuous.  The
    # embeddings of 512 molecules of an independent batch are used instead.
    with torch.no_grad():
        gi, _ = synthetic.make_batch(512, seed=79)
        state["output_regressor.x_tr"] = net_ref.representation(ref_graph(gi), state, CONFIG, layer="SAGEConv").clone()
    net.load_state_dict(state)

    def oracle(dtype):
        p = {k: v.detach().clone().to(dtype).requires_grad_(True) for k, v in state.items()}
        h = net_ref.representation(ref_graph(g), p, CONFIG, layer="SAGEConv")
        loss = gp_ref.vgp_loss(h, y.to(dtype), net_ref.vgp_params(p))
        loss.backward()
        with torch.no_grad():
            q = {k: v.detach() for k, v in p.items()}
            h_te = net_ref.representation(ref_graph(g_te), q, CONFIG, layer="SAGEConv")
            mean, cov = gp_ref.vgp_forward(h_te, net_ref.vgp_params(q))
        return {"loss": loss.detach(), "h": h.detach(), "grads": {k: v.grad for k, v in p.items()},
                "mean": mean.flatten(), "cov": cov, "var": torch.diagonal(cov).clone()}

    o32, o64 = oracle(torch.float32), oracle(torch.float64)
    net = net.to(DEV)
    gd, yd = g.to(DEV), y.to(DEV)
    close("embeddings h (element-wise)", net.representation(gd), o32["h"], o64["h"], 1e-4, elementwise=True)
    net.zero_grad()
    loss = net.loss(gd, yd)
    assert loss.dtype == torch.float64
    scalar_close("VGP loss", loss, o32["loss"], o64["loss"], 1e-3)
    loss.backward()
    for k, p in net.named_parameters():
        assert p.grad is not None and torch.isfinite(p.grad).all(), k
        close("grad " + k.replace("representation.", "rep.").replace("output_regressor.", "gp."), p.grad,
              o32["grads"][k], o64["grads"][k], 5e-3)
    with torch.no_grad():
        h_te = net.representation(g_te.to(DEV))
        mean, cov = net.output_regressor.forward(h_te)
        dist = net.condition(g_te.to(DEV))
    assert mean.shape == (96, 1)
    close("VGP predictive mean (element-wise)", mean, o32["mean"], o64["mean"], 1e-4, elementwise=True)
    close("VGP predictive covariance", cov, o32["cov"], o64["cov"], 1e-4)
    close("VGP predictive variance", dist.variance, o32["var"], o64["var"], 1e-4, elementwise=True)


@pytest.mark.parametrize("kind", ["rbf", "matern52"])
def test_blocked_sharded_sweep_at_n4600_matches_oracle(kind):
    """The column-block sweep that the multi-GPU path runs per rank (single process: every block is
    local) against the ORACLE, not only against the fused single-GPU path: N = 4 600, D = 128, blocks of
    512 (nine blocks, the last one ragged).  Gradient bar: 2e-4 for the embeddings and targets, 1e-3 for
    the three hyper-parameters -- their gradients are sums of ~2e7 terms of both signs, and the fp32
    oracle itself sits ~3e-4 away from the fp64 one (printed)."""
    gen = torch.Generator().manual_seed(7)
    n, d = 4600, 128
    x = torch.randn(n, d, generator=gen)
    y = torch.randn(n, 1, generator=gen)
    sc = (0.1 * torch.randn(d if kind == "rbf" else 1, generator=gen)) - 1.5
    if kind == "matern52":
        sc = sc.reshape(())
    var, ls = torch.tensor(0.2), torch.tensor(-2.0)
    okernel = "matern52_safe" if kind == "matern52" else "rbf"

    def oracle(dtype):
        ts = [t.clone().to(dtype).requires_grad_(True) for t in (x, y, sc, var, ls)]
        loss = gp_ref.exact_gp_loss(*ts, kernel=okernel)
        loss.sum().backward()
        return loss.detach(), [t.grad for t in ts]

    (l32, g32), (l64, g64) = oracle(torch.float32), oracle(torch.float64)
    ts = [t.clone().to(DEV).requires_grad_(True) for t in (x, y, sc, var, ls)]
    loss = sharded_gp.sharded_exact_gp_loss(*ts, kind=kind, block=512)
    loss.sum().backward()
    scalar_close("sharded NLL", loss.sum(), l32.sum(), l64.sum(), 1e-3)
    for name, t, a, b, tol in zip(["dx", "dy", "dscale", "dvariance", "dlog_sigma"], ts, g32, g64,
                                  [2e-4, 2e-4, 1e-3, 1e-3, 1e-3]):
        close("sharded " + name, t.grad, a, b, tol)


def test_train_once_closure_with_lbfgs_style_optimizer():
    """pinot/app/experiment.py:78-89 verbatim in shape: `optimizer.step(closure)` where the closure
    zeroes the gradients, sums `net.loss(g, y)` and calls backward -- with an optimiser that really
    re-evaluates the closure (LBFGS) and with Adam; the loss must go down and stay finite."""
    torch.manual_seed(0)
    g, y = synthetic.make_batch(256, seed=11, device=DEV)
    for make_opt in (lambda ps: torch.optim.LBFGS(ps, lr=0.1, max_iter=4),
                     lambda ps: torch.optim.Adam(ps, 1e-2)):
        net = make_net("GraphConv", ExactGaussianProcessRegressor).to(DEV)
        with torch.no_grad():
            net.output_regressor.kernel.scale.fill_(-3.0)
        optimizer = make_opt(list(net.parameters()))
        data = [(g, y)]
        losses = []
        for _ in range(3):
            for x in data:
                def l():
                    optimizer.zero_grad()
                    loss = torch.sum(net.loss(*x))
                    loss.backward()
                    losses.append(float(loss))
                    return loss
                optimizer.step(l)
        assert all(math.isfinite(v) for v in losses)
        assert losses[-1] < losses[0]
