"""Generate the golden fixtures in this directory by running the UNMODIFIED reference.

Run in the build container only (needs /root/reference):

    python tests/golden/make_golden.py

The reference's own ``pinot.Net`` / ``pinot.representation.Sequential`` / ``dgl_legacy.GN`` /
``pinot.regressors.*`` code executes on CPU; ``import dgl`` inside it resolves to the oracle's DGL
restatement (``oracle/dgl_ref.py``) because DGL itself cannot be installed.
Each fixture stores the inputs (graph arrays, features, targets), the ``state_dict`` and the
reference's outputs (embeddings, loss, every parameter gradient, predictive mean / covariance).
"""
import os
import sys
import warnings

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

warnings.simplefilter("ignore")

from oracle import dgl_ref, ref_loader  # noqa: E402
from pinot_b200 import synthetic  # noqa: E402  (data generator only: numpy/torch on CPU)

pinot = ref_loader.load()
from pinot.regressors import ExactGaussianProcessRegressor, VariationalGaussianProcessRegressor  # noqa: E402
from pinot.regressors.kernels.rbf import RBF  # noqa: E402
from pinot.regressors.kernels.matern import Matern52  # noqa: E402


def graph_arrays(g):
    """pinot_b200 BatchedGraph (CPU) -> plain arrays."""
    src, dst = g.edges()
    return {
        "src": src.clone(), "dst": dst.clone(), "num_nodes": g.number_of_nodes(),
        "batch_num_nodes": g.batch_num_nodes().clone(), "batch_num_edges": g.batch_num_edges().clone(),
        "h": g.ndata["h"].clone(),
    }


def to_ref_graph(a):
    return dgl_ref.from_arrays(a["src"], a["dst"], a["num_nodes"], a["batch_num_nodes"].tolist(),
                               a["batch_num_edges"].tolist(), ndata={"h": a["h"]})


def with_isolated_atom(g):
    """Append a one-atom molecule (no bonds): the zero-in-degree case (methane in ESOL)."""
    from pinot_b200.graph import BatchedGraph, batch
    single = BatchedGraph(num_nodes=1)
    feat = torch.zeros(1, 117)
    feat[0, 6] = 1.0
    feat[0, 100] = feat[0, 101] = 4.0
    feat[0, 105] = 12.011
    feat[0, 114] = 1.0
    single.ndata["h"] = feat
    return batch([g, single])


def grads_of(net):
    return {k: (p.grad.clone() if p.grad is not None else None) for k, p in net.named_parameters()}


def net_case(name, layer, config, n_tr, n_te, seed, head, kernel=None, head_kwargs=None, tweak=None,
             isolated=False, input_units=128):
    torch.manual_seed(seed)
    g_tr, y_tr = synthetic.make_batch(n_tr, seed=1000 + seed, mean_atoms=14.0, std_atoms=5.0)
    g_te, _ = synthetic.make_batch(n_te, seed=2000 + seed, mean_atoms=14.0, std_atoms=5.0)
    if isolated:
        g_tr = with_isolated_atom(g_tr)
        y_tr = torch.cat([y_tr, torch.tensor([[0.25]])])
    a_tr, a_te = graph_arrays(g_tr), graph_arrays(g_te)
    rg_tr, rg_te = to_ref_graph(a_tr), to_ref_graph(a_te)
    rep = pinot.representation.Sequential(pinot.representation.dgl_legacy.gn(layer), config, input_units=input_units)
    kw = dict(head_kwargs or {})
    if kernel is not None:
        kw["kernel"] = kernel
    net = pinot.Net(rep, output_regressor_class=head, **kw)
    if tweak is not None:
        with torch.no_grad():
            tweak(net)
    state = {k: v.clone() for k, v in net.state_dict().items()}
    h = net.representation(rg_tr).detach().clone()
    net.zero_grad()
    loss = net.loss(rg_tr, y_tr)
    loss.sum().backward()
    out = {
        "name": name, "layer": layer, "config": config, "input_units": input_units,
        "head": head.__name__, "kernel": "matern52" if kernel is not None and "atern" in getattr(kernel, "__name__", "matern") else "rbf",
        "head_kwargs": head_kwargs or {},
        "train": a_tr, "test": a_te, "y": y_tr, "state_dict": state,
        "h": h, "loss": loss.detach().clone(), "grads": grads_of(net),
    }
    with torch.no_grad():
        if head is ExactGaussianProcessRegressor:
            dist = net.condition(rg_te)
        else:
            dist = net.condition(rg_te)
        out["cond_mean"] = dist.mean.clone()
        out["cond_cov"] = dist.covariance_matrix.clone()
        out["cond_var"] = dist.variance.clone()
    torch.save(out, os.path.join(HERE, name + ".pt"))
    print("%-28s loss %-14s |h| %.3f  cov diag min %.3e" % (
        name, float(loss.sum()), float(h.norm()), float(out["cond_var"].min())))


def matern_unit(in_features):
    return Matern52()


def gp_case(name, head, n, n_te, d, seed, kernel_name="rbf", scale=-0.5, variance=0.3, log_sigma=-2.0, x_std=0.5,
            head_kwargs=None):
    torch.manual_seed(seed)
    x = (x_std * torch.randn(n, d)).requires_grad_(True)
    y = torch.randn(n, 1)
    x_te = x_std * torch.randn(n_te, d)
    kern = RBF if kernel_name == "rbf" else matern_unit
    reg = head(d, kernel=kern, **(head_kwargs or {}))
    with torch.no_grad():
        reg.kernel.scale.fill_(scale)
        reg.kernel.variance.fill_(variance)
        reg.log_sigma.fill_(log_sigma)
        if head is VariationalGaussianProcessRegressor:
            reg.y_tr_sigma_tril.copy_(0.05 * torch.randn_like(reg.y_tr_sigma_tril))
    state = {k: v.clone() for k, v in reg.state_dict().items()}
    loss = reg.loss(x, y)
    loss.sum().backward()
    out = {
        "name": name, "head": head.__name__, "kernel": kernel_name, "head_kwargs": head_kwargs or {},
        "x": x.detach().clone(), "y": y, "x_te": x_te, "state_dict": state,
        "loss": loss.detach().clone(), "grads": grads_of(reg), "grad_x": x.grad.clone(),
        "k_xx": reg.kernel(x.detach(), x.detach()).detach().clone()[:64, :64],
        "k_xte": reg.kernel(x.detach()[:64], x_te[:32]).detach().clone(),
    }
    with torch.no_grad():
        if head is ExactGaussianProcessRegressor:
            dist = reg.condition(x_te, x_tr=x.detach(), y_tr=y)
        else:
            dist = reg.condition(x_te)
        out["cond_mean"] = dist.mean.clone()
        out["cond_cov"] = dist.covariance_matrix.clone()
        out["cond_var"] = dist.variance.clone()
    torch.save(out, os.path.join(HERE, name + ".pt"))
    print("%-28s loss %-14s dtype %s" % (name, float(loss.sum()), loss.dtype))


def main():
    t3 = ["tanh"]
    net_case("net_gcn32_exact_rbf", "GraphConv", [32, "tanh", 32, "tanh", 32, "tanh"], 12, 6, 0,
             ExactGaussianProcessRegressor, isolated=True)

    def trained_like(net):
        net.output_regressor.kernel.scale.fill_(-3.0)
        net.output_regressor.kernel.variance.fill_(0.5)
        net.output_regressor.log_sigma.fill_(-2.0)

    net_case("net_gcn128_exact_rbf", "GraphConv", [128, "tanh", 128, "tanh", 128, "tanh"], 24, 8, 1,
             ExactGaussianProcessRegressor, tweak=trained_like)
    net_case("net_gcn64to32_exact_rbf", "GraphConv", [64, "tanh", 32, "tanh"], 16, 5, 2,
             ExactGaussianProcessRegressor, tweak=trained_like)

    def matern_like(net):
        net.output_regressor.kernel.scale.fill_(-2.5)
        net.output_regressor.log_sigma.fill_(-2.0)

    net_case("net_sage128_exact_matern", "SAGEConv", [128, "tanh", 64, "tanh"], 16, 5, 3,
             ExactGaussianProcessRegressor, kernel=matern_unit, tweak=matern_like)

    def vgp_like(net):
        net.output_regressor.kernel.scale.fill_(-2.0)
        r = net.output_regressor
        r.x_tr.copy_(3.0 * torch.randn_like(r.x_tr))
        r.y_tr_sigma_tril.copy_(0.05 * torch.randn_like(r.y_tr_sigma_tril))

    net_case("net_sage64_vgp", "SAGEConv", [64, "tanh", 64, "tanh"], 20, 6, 4,
             VariationalGaussianProcessRegressor, head_kwargs={"n_inducing_points": 24}, tweak=vgp_like)

    gp_case("gp_exact_rbf_n300", ExactGaussianProcessRegressor, 300, 50, 32, 10)
    gp_case("gp_exact_rbf_n131_d7", ExactGaussianProcessRegressor, 131, 17, 7, 11, scale=0.2)
    gp_case("gp_exact_matern_n150", ExactGaussianProcessRegressor, 150, 20, 16, 12, kernel_name="matern52", scale=-0.3)
    gp_case("gp_vgp_rbf_n200_m64", VariationalGaussianProcessRegressor, 200, 30, 32, 13,
            head_kwargs={"n_inducing_points": 64})
    _ = t3


if __name__ == "__main__":
    main()
