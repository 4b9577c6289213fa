"""BASELINE.json's full sizes, checked through size-independent properties (the CPU oracle would
take minutes to hours there): factor-times-transpose, solve residuals, linearity, adjointness."""
import pytest
import torch

import pinot_b200 as pb
from pinot_b200 import ops, synthetic
from conftest import rel_err

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.fixture(scope="module")
def c2():
    """configs[1]: 4.2k molecules, GraphConv 3x128 + exact GP."""
    torch.manual_seed(0)
    g, y = synthetic.make_batch(4200, seed=2666, device=DEV)
    net = pb.Net(pb.representation.Sequential(pb.representation.dgl_legacy.gn(), [128, "tanh", 128, "tanh", 128, "tanh"]),
                 output_regressor_class=pb.regressors.ExactGaussianProcessRegressor).to(DEV)
    with torch.no_grad():
        net.output_regressor.kernel.scale.fill_(-3.0)
    return net, g, y


def test_message_passing_is_linear_and_adjoint(c2):
    _, g, _ = c2
    csr = g.csr()
    n = g.number_of_nodes()
    x1, x2 = torch.randn(n, 128, device=DEV), torch.randn(n, 128, device=DEV)
    a = ops.aggregate(2.0 * x1 + x2, csr, "both")
    b = 2.0 * ops.aggregate(x1, csr, "both") + ops.aggregate(x2, csr, "both")
    assert rel_err(a, b) < 1e-6
    # <A x, y> == <x, A^T y>: forward kernel vs the transposed-CSR backward kernel
    xg = x1.clone().requires_grad_(True)
    ops.aggregate(xg, csr, "both").backward(x2)
    lhs = (ops.aggregate(x1, csr, "both").double() * x2.double()).sum()
    rhs = (x1.double() * xg.grad.double()).sum()
    assert abs(float(lhs - rhs)) < 1e-6 * abs(float(lhs))
    # sum readout conserves the total
    tot = ops.segment_readout(x1, csr).double().sum(0)
    assert rel_err(tot, x1.double().sum(0)) < 1e-6
    # "none" aggregation of ones counts in-degrees exactly
    deg = ops.aggregate(torch.ones(n, 4, device=DEV), csr, "none")[:, 0]
    assert torch.equal(deg.long(), g.in_degrees())


def test_cholesky_sweep_at_full_size(c2):
    net, g, y = c2
    with torch.no_grad():
        h = net.representation(g)
    n = h.shape[0]
    k = ops.gram("rbf", h, None, net.output_regressor.kernel.scale, net.output_regressor.kernel.variance,
                 diag_add=1e-5 + float(torch.exp(net.output_regressor.log_sigma)))
    assert torch.equal(k, k.t())
    rhs = torch.cat([y, torch.randn(n, 3, device=DEV)], dim=1)
    l, x = ops.chol_solve(k.detach(), rhs)
    assert torch.count_nonzero(torch.triu(l, 1)) == 0
    assert rel_err(l.double() @ l.double().t(), k.double()) < 1e-6          # L L^T = K
    assert rel_err(l.double() @ x.double(), rhs.double()) < 1e-5            # L X = B


def test_train_step_and_posterior_at_full_size(c2):
    net, g, y = c2
    net.zero_grad()
    loss = net.loss(g, y)
    loss.sum().backward()
    assert torch.isfinite(loss).all()
    for name, p in net.named_parameters():
        assert p.grad is not None and torch.isfinite(p.grad).all(), name
    # alpha solves (K + sigma^2 I) alpha = y: check through the predictive mean at the training
    # points, which must reproduce K alpha
    reg = net.output_regressor
    with torch.no_grad():
        h = net.representation(g)
        mean, var = ops.exact_gp_posterior(h[:512], h, y, reg.kernel.scale, reg.kernel.variance, reg.log_sigma,
                                           full_cov=False)
        assert torch.isfinite(mean).all() and torch.isfinite(var).all()
        assert float(var.min()) > -1e-4 and float(var.max()) <= float(torch.exp(reg.kernel.variance)) + 1e-4
        # interpolation property: residual at training points is sigma^2 * alpha (small for small sigma)
        k = ops.gram("rbf", h, None, reg.kernel.scale, reg.kernel.variance, diag_add=1e-5)
        kp = k + torch.exp(reg.log_sigma) * torch.eye(h.shape[0], device=DEV)
        alpha = torch.linalg.solve(kp.double(), y.double())
        assert rel_err(mean, ((k.double() + 0 * kp.double())[:512] @ alpha).flatten()) < 1e-3


def test_exact_gp_fit_at_n16k_properties():
    """configs[3] size: N = 16384 embeddings, H = 128 (X ~ 0.1 N(0,1), seed 0)."""
    torch.manual_seed(0)
    n, d = 16384, 128
    x = (0.1 * torch.randn(n, d)).to(DEV)
    y = torch.randn(n, 1).to(DEV)
    sc = torch.zeros(d, device=DEV)
    var = torch.zeros((), device=DEV)
    ls = torch.tensor(-3.0, device=DEV)
    xg = x.clone().requires_grad_(True)
    lsg = ls.clone().requires_grad_(True)
    loss = ops.exact_gp_loss(xg, y, sc, var, lsg)
    loss.sum().backward()
    assert torch.isfinite(loss).all() and torch.isfinite(xg.grad).all() and torch.isfinite(lsg.grad)
    # check the factor through a random probe: L (L^T v) = K v
    k = ops.gram("rbf", x, None, sc, var, diag_add=1e-5 + float(torch.exp(ls)))
    l = ops.cholesky(k)
    v = torch.randn(n, 1, device=DEV)
    lhs = ops.matmul(l, ops.matmul(l, v, True, False))
    assert rel_err(lhs, ops.matmul(k, v)) < 1e-5
