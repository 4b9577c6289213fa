"""pinot_b200.data.utils.batch / split against the reference's own functions (pinot/data/utils.py:207-282) run on
the oracle's DGL restatement, molecule by molecule: batch boundaries, node numbering, edge order, targets."""
import copy

import pytest
import torch

import pinot_b200 as pb
from pinot_b200 import synthetic
from oracle import dgl_ref, ref_loader


def _dataset(n, seed):
    """The same molecules as a list of single-graph entries, for both implementations."""
    g, y = synthetic.make_batch(n, seed=seed)
    ours, theirs = [], []
    for i, gi in enumerate(pb.unbatch(g)):
        s, d = gi.edges()
        gi.ndata["h"] = gi.ndata["h"].clone()
        ours.append((gi, y[i].clone()))
        rg = dgl_ref.from_arrays(s, d, gi.number_of_nodes(), ndata={"h": gi.ndata["h"].clone()})
        theirs.append((rg, y[i].clone()))
    return ours, theirs


def _same_batch(mine, ref):
    gm, ym = mine
    gr, yr = ref
    assert torch.equal(ym, yr) and ym.shape == yr.shape
    assert gm.number_of_nodes() == gr.number_of_nodes()
    assert gm.batch_num_nodes().tolist() == list(gr.node_counts())
    assert gm.batch_num_edges().tolist() == list(gr.edge_counts())
    sm, dm = gm.edges()
    assert sm.tolist() == [int(v) for v in gr.src] and dm.tolist() == [int(v) for v in gr.dst]
    assert torch.equal(gm.ndata["h"], gr.ndata["h"])


@pytest.mark.skipif(not ref_loader.available(), reason="reference tree not present")
@pytest.mark.parametrize("shuffle", [False, True])
@pytest.mark.parametrize("partial", [False, True])
def test_batch_matches_the_reference(shuffle, partial):
    ref = ref_loader.load()
    ours, theirs = _dataset(23, seed=3)
    mine = pb.data.utils.batch(ours, 5, seed=7, shuffle=shuffle, partial_batch=partial)
    want = ref.data.utils.batch(theirs, 5, seed=7, shuffle=shuffle, partial_batch=partial)
    assert len(mine) == len(want) == (5 if partial else 4)
    for a, b in zip(mine, want):
        _same_batch(a, b)


@pytest.mark.skipif(not ref_loader.available(), reason="reference tree not present")
def test_split_matches_the_reference():
    ref = ref_loader.load()
    ds = list(range(103))
    for partition in ([0.8, 0.1, 0.1], [1, 1], [3, 2, 2, 1], [0.5, 0.25]):
        assert pb.data.utils.split(ds, partition) == ref.data.utils.split(copy.copy(ds), partition)


def test_batch_semantics_without_the_reference():
    ours, _ = _dataset(11, seed=1)
    out = pb.data.utils.batch(ours, 4)
    assert [g.batch_size for g, _ in out] == [4, 4]                      # remainder of 3 dropped
    assert all(y.shape == (4, 1) for _, y in out)
    out = pb.data.utils.batch(ours, 4, partial_batch=True)
    assert [g.batch_size for g, _ in out] == [4, 4, 3]
    parts = pb.data.utils.split(list(range(10)), [0.8, 0.1, 0.1])
    assert [len(p) for p in parts] == [8, 1, 1]
    assert [len(p) for p in pb.data.utils.split(list(range(9)), [1, 1])] == [4, 4]   # truncation drops the last entry
