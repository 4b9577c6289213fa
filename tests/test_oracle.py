"""The oracle against the reference's own outputs (golden fixtures) and independent algebra."""
import pytest
import torch

from conftest import GP_EXACT_CASES, GP_VGP_CASES, NET_CASES, load_golden, rel_err
from oracle import dgl_ref, gp_ref, net_ref, ref_loader


def _ref_graph(a):
    return net_ref.graph_from_arrays(a["src"], a["dst"], a["num_nodes"], a["batch_num_nodes"].tolist(),
                                     a["batch_num_edges"].tolist(), a["h"])


def _leaf_params(state):
    return {k: v.clone().requires_grad_(True) for k, v in state.items()}


@pytest.mark.parametrize("name", GP_EXACT_CASES)
def test_gp_ref_exact_matches_reference(name):
    f = load_golden(name)
    p = _leaf_params(f["state_dict"])
    x = f["x"].clone().requires_grad_(True)
    loss = gp_ref.exact_gp_loss(x, f["y"], p["kernel.scale"], p["kernel.variance"], p["log_sigma"], kernel=f["kernel"])
    assert loss.shape == (1, 1)
    assert rel_err(loss, f["loss"]) < 1e-5
    loss.sum().backward()
    assert rel_err(x.grad, f["grad_x"], nan_equal=True) < 1e-3
    for k, g in f["grads"].items():
        assert rel_err(p[k].grad, g, nan_equal=True) < 1e-3, k
    with torch.no_grad():
        kern = gp_ref.KERNELS[f["kernel"]]
        k_xx = kern(f["x"], f["x"], f["state_dict"]["kernel.scale"], f["state_dict"]["kernel.variance"])[:64, :64]
        assert rel_err(k_xx, f["k_xx"]) < 1e-6
        mean, cov = gp_ref.exact_gp_condition(f["x_te"], f["x"], f["y"], f["state_dict"]["kernel.scale"],
                                              f["state_dict"]["kernel.variance"], f["state_dict"]["log_sigma"],
                                              kernel=f["kernel"])
    assert rel_err(mean, f["cond_mean"]) < 1e-4
    assert rel_err(cov, f["cond_cov"]) < 1e-4


@pytest.mark.parametrize("name", GP_VGP_CASES)
def test_gp_ref_vgp_matches_reference(name):
    f = load_golden(name)
    p = _leaf_params(f["state_dict"])
    x = f["x"].clone().requires_grad_(True)
    loss = gp_ref.vgp_loss(x, f["y"], p, kernel=f["kernel"])
    assert loss.dtype == torch.float64
    assert rel_err(loss, f["loss"]) < 1e-5
    loss.backward()
    assert rel_err(x.grad, f["grad_x"]) < 1e-3
    for k, g in f["grads"].items():
        assert rel_err(p[k].grad, g) < 2e-3, k
    with torch.no_grad():
        mean, cov = gp_ref.vgp_forward(f["x_te"], f["state_dict"], kernel=f["kernel"])
    assert rel_err(mean, f["cond_mean"]) < 1e-4
    assert rel_err(cov, f["cond_cov"]) < 1e-4


@pytest.mark.parametrize("name", NET_CASES)
def test_net_ref_matches_reference(name):
    f = load_golden(name)
    p = _leaf_params(f["state_dict"])
    g = _ref_graph(f["train"])
    if f["head"] == "ExactGaussianProcessRegressor":
        loss, h = net_ref.exact_net_loss(g, f["y"], p, f["config"], layer=f["layer"], kernel=f["kernel"])
    else:
        h = net_ref.representation(g, p, f["config"], f["layer"])
        loss = gp_ref.vgp_loss(h, f["y"], net_ref.vgp_params(p), kernel=f["kernel"])
    assert rel_err(h, f["h"]) < 1e-5
    assert rel_err(loss, f["loss"]) < 1e-5
    loss.sum().backward()
    for k, gr in f["grads"].items():
        if gr is None:
            continue
        assert rel_err(p[k].grad, gr, nan_equal=True) < 5e-3, k
    with torch.no_grad():
        if f["head"] == "ExactGaussianProcessRegressor":
            mean, cov = net_ref.exact_net_condition(_ref_graph(f["test"]), g, f["y"], f["state_dict"], f["config"],
                                                    layer=f["layer"], kernel=f["kernel"])
        else:
            h_te = net_ref.representation(_ref_graph(f["test"]), f["state_dict"], f["config"], f["layer"])
            mean, cov = gp_ref.vgp_forward(h_te, net_ref.vgp_params(f["state_dict"]), kernel=f["kernel"])
    assert rel_err(mean, f["cond_mean"]) < 1e-4
    assert rel_err(cov, f["cond_cov"]) < 1e-4


def test_dgl_ref_graphconv_equals_dense_adjacency_algebra():
    """GraphConv(norm='both') == D_in^-1/2 A D_out^-1/2 X W + b, SAGE mean == D_in^-1 A X (fp64)."""
    torch.manual_seed(0)
    f = load_golden("net_gcn32_exact_rbf")  # contains an isolated (zero-degree) atom
    g = _ref_graph(f["train"])
    n = g.number_of_nodes()
    a = g.dense_adjacency()
    x = torch.randn(n, 24, dtype=torch.float64)
    din = a.sum(1).clamp(min=1)
    dout = a.sum(0).clamp(min=1)
    assert (a.sum(1) == 0).any(), "fixture must exercise the zero-in-degree clamp"
    for fin, fout in [(24, 40), (24, 8)]:
        w = torch.randn(fin, fout, dtype=torch.float64)
        b = torch.randn(fout, dtype=torch.float64)
        want = (din.pow(-0.5)[:, None] * (a @ (dout.pow(-0.5)[:, None] * x))) @ w + b
        got = dgl_ref.graph_conv(g, x, w, b, "both")
        assert rel_err(got, want) < 1e-12
    ws, wn = torch.randn(16, 24, dtype=torch.float64), torch.randn(16, 24, dtype=torch.float64)
    bs, bn = torch.randn(16, dtype=torch.float64), torch.randn(16, dtype=torch.float64)
    want = x @ ws.t() + bs + ((a @ x) / din[:, None]) @ wn.t() + bn
    assert rel_err(dgl_ref.sage_conv_mean(g, x, ws, bs, wn, bn), want) < 1e-12
    # sum_nodes == block indicator matrix
    g.ndata["t"] = x
    seg = torch.zeros(g.batch_size, n, dtype=torch.float64)
    off = 0
    for i, c in enumerate(g.node_counts()):
        seg[i, off:off + c] = 1
        off += c
    assert rel_err(dgl_ref.sum_nodes(g, "t"), seg @ x) < 1e-12


def test_dgl_ref_against_networkx():
    """Second opinion on the adjacency/degree bookkeeping from an independent graph library."""
    nx = pytest.importorskip("networkx")
    f = load_golden("net_gcn64to32_exact_rbf")
    a = f["train"]
    g = _ref_graph(a)
    G = nx.MultiDiGraph()
    G.add_nodes_from(range(a["num_nodes"]))
    G.add_edges_from(zip(a["src"].tolist(), a["dst"].tolist()))
    assert [G.in_degree(i) for i in range(a["num_nodes"])] == g.in_degrees().tolist()
    assert [G.out_degree(i) for i in range(a["num_nodes"])] == g.out_degrees().tolist()
    A = torch.tensor(nx.to_numpy_array(G, nodelist=range(a["num_nodes"]))).t()  # A[i, j] = #edges j -> i
    assert torch.equal(A, g.dense_adjacency())
    # molecules are separate connected components, one per batched graph
    assert nx.number_weakly_connected_components(G) == g.batch_size


@pytest.mark.skipif(not ref_loader.available(), reason="reference tree not present (GPU box)")
def test_gp_ref_against_live_reference():
    """Fresh random problem, reference imported live: loss, gradients, posterior."""
    import warnings
    warnings.simplefilter("ignore")
    ref_loader.load()
    from pinot.regressors import ExactGaussianProcessRegressor
    torch.manual_seed(5)
    x = (0.7 * torch.randn(90, 12)).requires_grad_(True)
    y = torch.randn(90, 1)
    x_te = 0.7 * torch.randn(11, 12)
    reg = ExactGaussianProcessRegressor(12)
    with torch.no_grad():
        reg.kernel.scale.fill_(-0.4)
    want = reg.loss(x, y)
    want.sum().backward()
    gx = x.grad.clone()
    x2 = x.detach().clone().requires_grad_(True)
    got = gp_ref.exact_gp_loss(x2, y, reg.kernel.scale.detach(), reg.kernel.variance.detach(), reg.log_sigma.detach())
    got.sum().backward()
    assert rel_err(got, want) < 1e-6
    assert rel_err(x2.grad, gx) < 1e-3
    with torch.no_grad():
        d = reg.condition(x_te, x_tr=x.detach(), y_tr=y)
        mean, cov = gp_ref.exact_gp_condition(x_te, x.detach(), y, reg.kernel.scale, reg.kernel.variance, reg.log_sigma)
    assert rel_err(mean, d.mean) < 1e-5
    assert rel_err(cov, d.covariance_matrix) < 1e-5
