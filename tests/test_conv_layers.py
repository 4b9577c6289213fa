"""GINConv / SGConv / TAGConv (the other layer types pinot's SequentialMix and gn()
name that reduce to the CSR aggregate): the oracle restatements against dense-adjacency algebra, and the
pinot_b200 layers against the oracle -- on the CPU through the C-ABI stand-ins (layer wiring, autograd
plumbing) and on the GPU through the kernels."""
import pytest
import torch

import pinot_b200 as pb
from pinot_b200 import synthetic
from pinot_b200.dgl.nn.pytorch import conv
from conftest import rel_err
from oracle import dgl_ref, net_ref
import cabi_standin


def _graphs(n_mol=12, seed=21):
    g, _ = synthetic.make_batch(n_mol, seed=seed, mean_atoms=9.0, std_atoms=3.0)
    s, d = g.edges()
    rg = net_ref.graph_from_arrays(s, d, g.number_of_nodes(), g.batch_num_nodes().tolist(),
                                   g.batch_num_edges().tolist(), g.ndata["h"])
    return g, rg


def _dense(rg, dtype=torch.float64):
    a = torch.zeros(rg.n, rg.n, dtype=dtype)
    for u, v in zip(rg.src, rg.dst):
        a[v, u] += 1.0          # row = destination
    return a


def test_oracle_layers_equal_dense_adjacency_algebra():
    _, rg = _graphs()
    torch.manual_seed(0)
    x = torch.randn(rg.n, 6, dtype=torch.float64)
    a = _dense(rg)
    deg = a.sum(1).clamp(min=1)
    s = torch.diag(deg.pow(-0.5)) @ a @ torch.diag(deg.pow(-0.5))
    assert rel_err(dgl_ref.gin_conv(rg, x, eps=0.3), 1.3 * x + a @ x) < 1e-12
    w = torch.randn(5, 6, dtype=torch.float64)
    b = torch.randn(5, dtype=torch.float64)
    assert rel_err(dgl_ref.sg_conv(rg, x, w, b, k=3), (s @ s @ s @ x) @ w.t() + b) < 1e-12
    w2 = torch.randn(5, 18, dtype=torch.float64)
    assert rel_err(dgl_ref.tag_conv(rg, x, w2, b, k=2), torch.cat([x, s @ x, s @ s @ x], 1) @ w2.t() + b) < 1e-12


def _layer_pairs(fin, fout):
    torch.manual_seed(1)
    pairs = [
        (conv.GINConv(torch.nn.Linear(fin, fout), "sum"), dgl_ref.GINConv(torch.nn.Linear(fin, fout), "sum")),
        (conv.GINConv(None, "mean", init_eps=0.25), dgl_ref.GINConv(None, "mean", init_eps=0.25)),
        (conv.SGConv(fin, fout, k=2), dgl_ref.SGConv(fin, fout, k=2)),
        (conv.TAGConv(fin, fout, k=2), dgl_ref.TAGConv(fin, fout, k=2)),
        (conv.TAGConv(fin, fout, k=3, bias=False), dgl_ref.TAGConv(fin, fout, k=3, bias=False)),
    ]
    for ours, ref in pairs:
        assert list(ours.state_dict().keys()) == list(ref.state_dict().keys())
        ours.load_state_dict(ref.state_dict())
    return pairs


def _check_layers(device):
    g, rg = _graphs()
    fin, fout = 16, 8
    x = torch.randn(g.number_of_nodes(), fin)
    for ours, ref in _layer_pairs(fin, fout):
        xr = x.clone().requires_grad_(True)
        want = ref(rg, xr)
        want.pow(2).sum().backward()
        ours = ours.to(device)
        xo = x.clone().to(device).requires_grad_(True)
        got = ours(g.to(device), xo)
        got.pow(2).sum().backward()
        name = type(ours).__name__
        assert rel_err(got, want) < 1e-5, name
        assert rel_err(xo.grad, xr.grad) < 1e-5, name
        for (k, po), (_, pr) in zip(ours.named_parameters(), ref.named_parameters()):
            assert rel_err(po.grad, pr.grad) < 1e-5, (name, k)
        # fused activation hand-off used by Sequential
        got_t = ours(g.to(device), xo.detach(), _act="tanh")
        assert rel_err(got_t, torch.tanh(want.detach())) < 1e-5, name


def test_layers_match_oracle_through_cpu_standins():
    with cabi_standin.installed():
        _check_layers("cpu")


@pytest.mark.gpu
def test_layers_match_oracle_on_gpu():
    _check_layers("cuda")


def _check_factories(device):
    g, rg = _graphs()
    # gn("GINConv") hands back the bare layer with a Linear apply function (dgl_legacy.py:60-64)
    rep = pb.representation.Sequential(pb.representation.dgl_legacy.gn("GINConv"), [8, "tanh", 8, "tanh"])
    assert isinstance(rep.d0, conv.GINConv) and rep.d0.apply_func.out_features == 8
    net = pb.Net(rep, output_regressor_class=pb.regressors.ExactGaussianProcessRegressor)
    assert net.representation_out_features == 8
    rep2 = pb.representation.Sequential(pb.representation.dgl_legacy.gn("TAGConv"), [8, "tanh"])
    assert rep2.d0.gn._k == 2 and rep2.d0.gn.lin.in_features == 128 * 3       # DEFAULT_MODEL_KWARGS
    mix = pb.representation.SequentialMix(["GraphConv", "8", "activation", "tanh", "TAGConv", "8", "2", "SGConv", "4", "3",
                                           "GINConv", "4", "dropout", "0.0"])
    assert mix.exes == ["conv0", "act1", "conv2", "conv3", "conv4", "drop5"]
    h = mix.to(device)(g.to(device))
    assert h.shape == (g.batch_size, 4) and torch.isfinite(h).all()
    with pytest.raises(NotImplementedError):
        pb.representation.SequentialMix(["GATConv", "8", "3"])
    return rep, rg, g


def test_factories_and_sequential_mix_through_cpu_standins():
    with cabi_standin.installed():
        rep, rg, g = _check_factories("cpu")
        # Sequential with GIN layers equals the oracle composition
        h = rep(g)
        x = torch.tanh(g.ndata["h"] @ rep.f_in[0].weight.t() + rep.f_in[0].bias)
        for name in ("d0", "d2"):
            lay = getattr(rep, name)
            x = torch.tanh(dgl_ref.gin_conv(rg, x, 0.0) @ lay.apply_func.weight.t() + lay.apply_func.bias)
        with rg.local_scope():
            rg.ndata["h"] = x
            want = dgl_ref.sum_nodes(rg, "h")
        assert rel_err(h, want) < 1e-5


@pytest.mark.gpu
def test_factories_and_sequential_mix_on_gpu():
    _check_factories("cuda")
