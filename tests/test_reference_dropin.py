"""The UNMODIFIED reference's loops on top of the ``dgl`` facade.

``pinot_b200.dgl.install_as_dgl()`` registers the facade as ``dgl``; the reference package is then imported
from /root/reference as it is and its own ``pinot.Net`` / ``pinot.representation.Sequential`` /
``pinot.representation.dgl_legacy.gn`` / regressors / ``pinot.app.experiment.Train``
(pinot/app/experiment.py:78-89) / ``pinot.active.experiment.BayesOptExperiment.run``
(pinot/active/experiment.py:361-402) / ``pinot.active.acquisition.thompson_sampling`` run unchanged: every
``dgl.*`` call they make (DGLGraph, batch, unbatch, sum_nodes, local_scope, ndata, batch_size,
dgl.nn.pytorch.conv.GraphConv) lands in pinot_b200.  There is no GPU here, so the C-ABI entry points behind
the facade are replaced by tests/cabi_standin.py; what this test pins is the host-side contract.  Skipped
where /root/reference is absent (the GPU box).
"""
import importlib.machinery
import sys
import types

import pytest
import torch

import pinot_b200 as pb
from pinot_b200 import synthetic
from oracle import dgl_ref, ref_loader
import cabi_standin

pytestmark = pytest.mark.skipif(not ref_loader.available(), reason="reference tree not present")


def _stub(name, **attrs):
    m = types.ModuleType(name)
    m.__spec__ = importlib.machinery.ModuleSpec(name, None)
    m.__path__ = []
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


@pytest.fixture()
def reference_on_facade():
    """Import the reference with ``dgl`` = the pinot_b200 facade; restore sys.modules afterwards so that
    other tests (which import the reference on the oracle's restatement) are not affected."""
    saved = {k: v for k, v in sys.modules.items()
             if k == "pinot" or k.startswith("pinot.") or k == "dgl" or k.startswith("dgl.")
             or k.startswith("rdkit") or k.startswith("matplotlib")}
    for k in saved:
        del sys.modules[k]
    pb.dgl.install_as_dgl(force=True)
    if "rdkit" not in sys.modules:
        _stub("rdkit", Chem=object())
        _stub("rdkit.Chem")
    if "matplotlib" not in sys.modules:
        _stub("matplotlib").pyplot = _stub("matplotlib.pyplot")
    if ref_loader.REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, ref_loader.REFERENCE_ROOT)
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        import pinot
    assert pinot.__file__.startswith(ref_loader.REFERENCE_ROOT)
    import dgl
    assert dgl is pb.dgl
    try:
        with cabi_standin.installed():
            yield pinot
    finally:
        for k in [k for k in sys.modules if k == "pinot" or k.startswith("pinot.") or k == "dgl" or k.startswith("dgl.")]:
            del sys.modules[k]
        sys.modules.update(saved)


def _reference_net(pinot, width=16):
    rep = pinot.representation.Sequential(pinot.representation.dgl_legacy.gn(), [width, "tanh", width, "tanh"])
    return pinot.Net(rep, output_regressor_class=pinot.regressors.ExactGaussianProcessRegressor)


def _oracle_loss(net, g, y, config):
    """The same weights through the oracle's DGL restatement + GP restatement."""
    from oracle import net_ref
    s, d = g.edges()
    rg = net_ref.graph_from_arrays(s, d, g.number_of_nodes(), g.batch_num_nodes().tolist(),
                                   g.batch_num_edges().tolist(), g.ndata["h"])
    p = {k: v.detach().clone() for k, v in net.state_dict().items()}
    return net_ref.exact_net_loss(rg, y, p, config)[0]


def test_reference_train_loop_runs_on_the_facade(reference_on_facade):
    pinot = reference_on_facade
    torch.manual_seed(0)
    g, y = synthetic.make_batch(24, seed=3, mean_atoms=10.0, std_atoms=3.0)
    net = _reference_net(pinot)
    # the facade's GraphConv really is the layer inside the reference's GN wrapper
    assert type(net.representation.d0.gn) is pb.dgl.nn.pytorch.conv.GraphConv
    with torch.no_grad():
        net.output_regressor.kernel.scale.fill_(-2.0)
    before = float(net.loss(g, y).sum())
    want = float(_oracle_loss(net, g, y, [16, "tanh", 16, "tanh"]).sum())
    assert abs(before - want) < 1e-3 * max(1.0, abs(want))
    trainer = pinot.app.experiment.Train(net=net, data=[(g, y)], optimizer=torch.optim.Adam(net.parameters(), 1e-2),
                                         n_epochs=2, record_interval=1)
    trainer.train()                                   # pinot/app/experiment.py:78-112, unmodified
    assert set(trainer.states.keys()) == {0, 1, "final"}
    after = float(net.loss(g, y).sum())
    assert after < before
    dist = net.condition(synthetic.make_batch(5, seed=4, mean_atoms=10.0, std_atoms=3.0)[0])
    assert dist.mean.shape == (5,) and torch.isfinite(dist.variance).all()


def test_reference_bayesopt_loop_runs_on_the_facade(reference_on_facade):
    pinot = reference_on_facade
    from pinot.active.experiment import BayesOptExperiment, _slice_fn_tuple, _collate_fn_graph
    torch.manual_seed(0)
    g, y = synthetic.make_batch(20, seed=5, mean_atoms=9.0, std_atoms=2.0)
    net = _reference_net(pinot, width=8)
    bo = BayesOptExperiment(
        net=net, data=(g, y), acquisition=pinot.active.acquisition.thompson_sampling,
        optimizer=lambda n: torch.optim.Adam(n.parameters(), 1e-2), num_epochs=2, q=3,
        slice_fn=_slice_fn_tuple, collate_fn=_collate_fn_graph, early_stopping=False)
    acquisitions = bo.run(num_rounds=2, seed=2666)    # pinot/active/experiment.py:361-402, unmodified
    assert len(acquisitions) == 2
    seen, unseen = acquisitions[-1]
    assert len(seen) == 1 + 3 and len(unseen) == 20 - 4 and not set(seen) & set(unseen)
    assert len(bo.seen) == 1 + 3 + 3
    # the re-batched graphs are genuine batched graphs of the facade with the right molecules
    gs, ys = bo.seen_data
    assert gs.batch_size == len(bo.seen) and ys.shape[0] == len(bo.seen)
    want = dgl_ref.batch([dgl_ref.unbatch(_to_ref(g))[i] for i in bo.seen])
    assert gs.edges()[0].tolist() == want.src and gs.edges()[1].tolist() == want.dst


def _to_ref(g):
    from oracle import net_ref
    s, d = g.edges()
    return net_ref.graph_from_arrays(s, d, g.number_of_nodes(), g.batch_num_nodes().tolist(),
                                     g.batch_num_edges().tolist(), g.ndata["h"])
