"""End-to-end parity with the reference: the golden fixtures hold outputs of the reference's own
Net / Sequential / regressors (tests/golden/make_golden.py); the same state_dict is loaded into
pinot_b200 and everything runs through the C-ABI kernels on the GPU.

Bars (north_star): embeddings and predictive mean / variance within 1e-4 relative in fp32, NLL
within 1e-3; gradients (not named by north_star) within 5e-3 of the reference's fp32 autograd."""
import pytest
import torch

import pinot_b200 as pb
from pinot_b200.graph import BatchedGraph
from conftest import GP_EXACT_CASES, GP_VGP_CASES, NET_CASES, load_golden, max_rel_err, rel_err
from test_api import build_net
from oracle import gp_ref

pytestmark = pytest.mark.gpu
DEV = "cuda"


def to_graph(a):
    g = BatchedGraph(a["src"], a["dst"], a["num_nodes"], a["batch_num_nodes"].tolist(), a["batch_num_edges"].tolist())
    g.ndata["h"] = a["h"]
    return g.to(DEV)


@pytest.mark.parametrize("name", NET_CASES)
def test_net_matches_reference(name):
    f = load_golden(name)
    net = build_net(f)
    net.load_state_dict(f["state_dict"])
    net = net.to(DEV)
    g, y = to_graph(f["train"]), f["y"].to(DEV)
    h = net.representation(g)
    assert h.shape == f["h"].shape
    # "within 1e-4 relative": norm-wise 1e-5, and element-wise 1e-4 for every entry that is not a
    # near-total cancellation of the sum readout (entries below 1% of the largest are measured
    # against that 1% floor)
    assert rel_err(h, f["h"]) < 1e-5
    assert max_rel_err(h, f["h"], floor=1e-2) < 1e-4, "embeddings must match within 1e-4 relative"
    net.zero_grad()
    loss = net.loss(g, y)
    assert loss.dtype == f["loss"].dtype and loss.shape == f["loss"].shape
    assert abs(float(loss.sum()) - float(f["loss"].sum())) < 1e-3 * max(1.0, abs(float(f["loss"].sum())))
    loss.sum().backward()
    params = dict(net.named_parameters())
    for k, gr in f["grads"].items():
        if gr is None or torch.isnan(gr).any():
            # the reference's Matern52 gradients are NaN (0 * inf at zero distance); ours must be finite
            if gr is not None:
                assert torch.isfinite(params[k].grad).all(), k
            continue
        scale = float(gr.abs().max())
        if scale < 1e-7:
            assert float(params[k].grad.abs().max()) < 1e-5, k
            continue
        assert rel_err(params[k].grad, gr) < 5e-3, k
    with torch.no_grad():
        dist = net.condition(to_graph(f["test"]))
    assert isinstance(dist, torch.distributions.Distribution)
    assert max_rel_err(dist.mean, f["cond_mean"], floor=1e-2) < 1e-3
    assert rel_err(dist.mean, f["cond_mean"]) < 1e-4
    assert max_rel_err(dist.variance, f["cond_var"], floor=1e-3) < 1e-3
    assert rel_err(dist.variance, f["cond_var"]) < 1e-4
    assert rel_err(dist.covariance_matrix, f["cond_cov"]) < 1e-4
    # consumers: _independent (pinot/metrics.py:13-27) and sampling (pinot/active/acquisition.py:101)
    assert dist.sample((3,)).shape == (3, f["cond_mean"].shape[0])
    assert torch.isfinite(dist.log_prob(f["cond_mean"].to(DEV)))
    if f["head"] == "ExactGaussianProcessRegressor":
        marg = net.condition_marginal(to_graph(f["test"]))
        assert rel_err(marg.mean, f["cond_mean"]) < 1e-4
        assert rel_err(marg.variance, f["cond_var"]) < 1e-4


@pytest.mark.parametrize("name", GP_EXACT_CASES)
def test_exact_gp_head_matches_reference(name):
    f = load_golden(name)
    kern = pb.regressors.kernels.RBF if f["kernel"] == "rbf" else (lambda d: pb.regressors.kernels.Matern52())
    reg = pb.regressors.ExactGaussianProcessRegressor(f["x"].shape[1], kernel=kern)
    reg.load_state_dict(f["state_dict"])
    reg = reg.to(DEV)
    x = f["x"].to(DEV).requires_grad_(True)
    y = f["y"].to(DEV)
    # the kernel module on its own (kernels/rbf.py:35-55)
    k = reg.kernel(x.detach(), x.detach())
    assert max_rel_err(k[:64, :64], f["k_xx"]) < 1e-4
    assert max_rel_err(reg.kernel(x.detach()[:64], f["x_te"][:32].to(DEV)), f["k_xte"]) < 1e-4
    loss = reg.loss(x, y)
    assert loss.shape == (1, 1)
    assert abs(float(loss) - float(f["loss"])) < 1e-3 * max(1.0, abs(float(f["loss"])))
    loss.sum().backward()
    if not torch.isnan(f["grad_x"]).any():
        assert rel_err(x.grad, f["grad_x"]) < 5e-3
        for kname, gr in f["grads"].items():
            assert rel_err(dict(reg.named_parameters())[kname].grad, gr) < 5e-3, kname
    else:
        assert torch.isfinite(x.grad).all()
        # finite-gradient oracle with identical forward values
        p = {kk: v.double().requires_grad_(True) for kk, v in f["state_dict"].items()}
        xr = f["x"].double().requires_grad_(True)
        gp_ref.exact_gp_loss(xr, f["y"].double(), p["kernel.scale"], p["kernel.variance"], p["log_sigma"],
                             kernel="matern52_safe").sum().backward()
        assert rel_err(x.grad, xr.grad) < 5e-3
        assert rel_err(reg.kernel.scale.grad, p["kernel.scale"].grad) < 5e-3
    with torch.no_grad():
        dist = reg.condition(f["x_te"].to(DEV), x_tr=x.detach(), y_tr=y)
    assert rel_err(dist.mean, f["cond_mean"]) < 1e-4
    assert rel_err(dist.variance, f["cond_var"]) < 1e-4
    assert rel_err(dist.covariance_matrix, f["cond_cov"]) < 1e-4
    # the unfused API-parity path (:85-121)
    with torch.no_grad():
        ktt, _, kxt, _, l_low, alpha = reg._get_kernel_and_auxiliary_variables(x.detach(), y, f["x_te"].to(DEV))
        assert rel_err(kxt @ alpha, f["cond_mean"]) < 1e-3


@pytest.mark.parametrize("name", GP_VGP_CASES)
def test_variational_gp_head_matches_reference(name):
    f = load_golden(name)
    reg = pb.regressors.VariationalGaussianProcessRegressor(f["x"].shape[1], **f["head_kwargs"])
    reg.load_state_dict(f["state_dict"])
    reg = reg.to(DEV)
    x = f["x"].to(DEV).requires_grad_(True)
    loss = reg.loss(x, f["y"].to(DEV))
    assert loss.dtype == torch.float64
    assert abs(float(loss) - float(f["loss"])) < 1e-3 * max(1.0, abs(float(f["loss"])))
    loss.backward()
    assert rel_err(x.grad, f["grad_x"]) < 5e-3
    for kname, gr in f["grads"].items():
        assert rel_err(dict(reg.named_parameters())[kname].grad, gr) < 1e-2, kname
    with torch.no_grad():
        mean, cov = reg.forward(f["x_te"].to(DEV))
        dist = reg.condition(f["x_te"].to(DEV))
    assert mean.shape == (f["x_te"].shape[0], 1)
    assert rel_err(mean, f["cond_mean"]) < 1e-4 and rel_err(cov, f["cond_cov"]) < 1e-4
    assert rel_err(dist.variance, f["cond_var"]) < 1e-4
