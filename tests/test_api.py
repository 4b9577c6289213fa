"""Drop-in surface: parameter names / shapes equal the reference's state_dict (golden fixtures),
constructor semantics of Net / Sequential / GN / regressors (host logic, CPU)."""
import copy

import pytest
import torch

import pinot_b200 as pb
from pinot_b200.regressors import (ExactGaussianProcessRegressor, NeuralNetworkRegressor,
                                   VariationalGaussianProcessRegressor)
from pinot_b200.regressors.kernels import RBF, Matern52
from conftest import NET_CASES, load_golden


def build_net(f):
    head = {"ExactGaussianProcessRegressor": ExactGaussianProcessRegressor,
            "VariationalGaussianProcessRegressor": VariationalGaussianProcessRegressor}[f["head"]]
    kw = dict(f["head_kwargs"])
    if f["kernel"] == "matern52":
        kw["kernel"] = lambda in_features: Matern52()
    rep = pb.representation.Sequential(pb.representation.dgl_legacy.gn(f["layer"]), f["config"],
                                       input_units=f["input_units"])
    return pb.Net(rep, output_regressor_class=head, **kw)


@pytest.mark.parametrize("name", NET_CASES)
def test_state_dict_is_interchangeable_with_the_reference(name):
    f = load_golden(name)
    net = build_net(f)
    ours = net.state_dict()
    assert list(ours.keys()) == list(f["state_dict"].keys())
    for k, v in f["state_dict"].items():
        assert ours[k].shape == v.shape and ours[k].dtype == v.dtype, k
    net.load_state_dict(f["state_dict"])  # strict
    assert net.has_exact_gp == (f["head"] == "ExactGaussianProcessRegressor")
    assert net.representation_out_features == [e for e in f["config"] if isinstance(e, int)][-1]
    copy.deepcopy(net)  # Train keeps deep copies of state (pinot/app/experiment.py:109-112)


def test_sequential_config_grammar():
    rep = pb.representation.Sequential(pb.representation.dgl_legacy.gn(), [32, "tanh", 0.1, "64", "relu"])
    assert rep.exes == ["d0", "a1", "o2", "d3", "a4"]
    assert rep.d0.in_features == 128 and rep.d0.out_features == 32 and rep.d3.out_features == 64
    assert rep.a1 is torch.tanh and isinstance(rep.o2, torch.nn.Dropout)
    assert rep.f_in[0].in_features == 117 and rep.f_in[0].out_features == 128
    assert rep.d0.gn.weight.shape == (128, 32)       # DGL GraphConv weight is (in, out)
    sage = pb.representation.dgl_legacy.gn("SAGEConv")(16, 8)
    assert sage.gn.fc_self.weight.shape == (8, 16) and sage.gn._aggre_type == "mean"
    with pytest.raises(NotImplementedError):
        pb.representation.dgl_legacy.gn("GATConv")(4, 4)


def test_net_accepts_any_module_with_out_features():
    """pinot/active/tests use Linear(1, 50) + Tanh as the representation."""
    rep = torch.nn.Sequential(torch.nn.Linear(1, 50), torch.nn.Tanh())
    net = pb.Net(rep, output_regressor_class=ExactGaussianProcessRegressor)
    assert net.representation_out_features == 50 and net.output_regressor.kernel.scale.shape == (50,)
    net2 = pb.Net(rep)  # default head is the neural-network regressor
    assert isinstance(net2.output_regressor, NeuralNetworkRegressor) and not net2.has_exact_gp
    d = net2.condition(torch.randn(7, 1))
    assert d.mean.shape == (7, 1)
    assert net2.loss(torch.randn(7, 1), torch.randn(7, 1)).dim() == 0


def test_regressor_constructors_match_reference_defaults():
    e = ExactGaussianProcessRegressor(16)
    assert float(e.log_sigma) == -3.0 and e.epsilon == 1e-5 and isinstance(e.kernel, RBF)
    assert e.kernel.scale.shape == (16,) and e.kernel.variance.shape == ()
    v = VariationalGaussianProcessRegressor(16)
    assert v.n_inducing_points == 100 and v.kl_loss_scaling == 1e-2
    assert v.x_tr.shape == (100, 16) and v.y_tr_mu.shape == (100, 1)
    assert v.y_tr_sigma_diag.shape == (100,) and v.y_tr_sigma_tril.shape == (4950,)
    assert float(v.x_tr.abs().max()) <= 1.0 and torch.all(v.y_tr_sigma_diag == -2.0)
    s = v._y_tr_sigma()
    assert torch.equal(s, torch.tril(s)) and torch.allclose(torch.diagonal(s), torch.exp(v.y_tr_sigma_diag))
    m = Matern52()
    assert m.scale.shape == () and RBF(5, ard=False).scale.shape == ()
    # quirk kept from the reference: kernel(in_features) hands in_features to Matern52's `scale`
    assert float(ExactGaussianProcessRegressor(7, kernel=Matern52).kernel.scale) == 7.0
    k = e._perturb(torch.zeros(3, 5))
    assert torch.equal(k, 1e-5 * torch.eye(3, 5))


def test_masked_scatter_fill_order_of_variational_scale():
    v = VariationalGaussianProcessRegressor(4, n_inducing_points=4)
    with torch.no_grad():
        v.y_tr_sigma_tril.copy_(torch.arange(1.0, 7.0))
    s = v._y_tr_sigma()
    assert s[1, 0] == 1 and s[2, 0] == 2 and s[2, 1] == 3 and s[3, 0] == 4 and s[3, 2] == 6  # row-major


def test_dgl_facade_install():
    import sys
    mod = pb.dgl.install_as_dgl(force=True)
    import dgl
    assert dgl is mod and dgl.nn.pytorch.conv.GraphConv is pb.dgl.nn.pytorch.conv.GraphConv
    assert hasattr(dgl, "batch") and hasattr(dgl, "unbatch") and hasattr(dgl, "sum_nodes") and hasattr(dgl, "DGLGraph")
    for k in ["dgl", "dgl.nn", "dgl.nn.pytorch", "dgl.nn.pytorch.conv"]:
        sys.modules.pop(k, None)
