"""Bit-exact batching / indexing against the oracle's DGL restatement (host logic, CPU)."""
import pytest
import torch

import pinot_b200 as pb
from pinot_b200 import synthetic
from pinot_b200.graph import BatchedGraph
from oracle import dgl_ref


def _to_ref(g):
    s, d = g.edges()
    r = dgl_ref.DGLGraph()
    r.add_nodes(g.number_of_nodes())
    r.add_edges(s, d)
    r.ndata["h"] = g.ndata["h"]
    return r


def test_batch_is_bit_exact_with_oracle_batch():
    mols, _ = synthetic.make_molecules(37, seed=3)
    got = pb.batch(mols)
    want = dgl_ref.batch([_to_ref(m) for m in mols])
    s, d = got.edges()
    assert s.tolist() == want.src and d.tolist() == want.dst
    assert got.batch_num_nodes().tolist() == want.batch_nodes
    assert got.batch_num_edges().tolist() == want.batch_edges
    assert torch.equal(got.ndata["h"], want.ndata["h"])
    assert got.batch_size == want.batch_size == 37
    # make_batch (vectorised generator) produces the very same batched graph
    direct, _ = synthetic.make_batch(37, seed=3)
    assert torch.equal(direct.edges()[0], s) and torch.equal(direct.edges()[1], d)


def test_unbatch_roundtrip_and_rebatch_of_a_selection():
    g, _ = synthetic.make_batch(20, seed=4)
    parts = pb.unbatch(g)
    assert len(parts) == 20
    again = pb.batch(parts)
    assert torch.equal(again.edges()[0], g.edges()[0]) and torch.equal(again.edges()[1], g.edges()[1])
    # the BO loop re-batches arbitrary selections (pinot/active/experiment.py:99-104)
    idxs = [7, 2, 19, 2]
    sel = pb.batch([parts[i] for i in idxs])
    ref_parts = dgl_ref.unbatch(dgl_ref.from_arrays(*g.edges(), g.number_of_nodes(), g.batch_num_nodes().tolist(),
                                                    g.batch_num_edges().tolist(), ndata={"h": g.ndata["h"]}))
    want = dgl_ref.batch([ref_parts[i] for i in idxs])
    assert sel.edges()[0].tolist() == want.src and sel.edges()[1].tolist() == want.dst
    assert torch.equal(sel.ndata["h"], want.ndata["h"])


def test_edge_layout_follows_from_rdkit_mol():
    """all bonds u->v, then all bonds v->u (pinot/graph.py:238-239)."""
    mols, _ = synthetic.make_molecules(10, seed=5)
    for m in mols:
        s, d = m.edges()
        nb = s.numel() // 2
        assert torch.equal(s[:nb], d[nb:]) and torch.equal(d[:nb], s[nb:])
        assert m.ndata["h"].shape == (m.number_of_nodes(), 117)
        assert torch.all(m.ndata["h"][:, :100].sum(1) == 1)
        assert int(m.in_degrees().max()) <= 4


def test_csr_views():
    g, _ = synthetic.make_batch(15, seed=6)
    c = g.csr()
    s, d = g.edges()
    n = g.number_of_nodes()
    ind = g.in_degrees()
    assert torch.equal(c.indptr[1:].long() - c.indptr[:-1].long(), ind)
    # row i lists the sources of the edges into i, in edge order
    for i in [0, 1, n // 2, n - 1]:
        want = s[d == i].tolist()
        assert c.indices[c.indptr[i]:c.indptr[i + 1]].tolist() == want
        want_t = d[s == i].tolist()
        assert c.indices_t[c.indptr_t[i]:c.indptr_t[i + 1]].tolist() == want_t
    assert torch.allclose(c.in_norm, ind.float().clamp(min=1).pow(-0.5))
    assert c.seg_offsets.tolist()[-1] == n and c.node_seg.numel() == n
    assert c.node_seg[c.seg_offsets[3]:c.seg_offsets[4]].unique().tolist() == [3]


def test_empty_and_degenerate_graphs():
    single = BatchedGraph(num_nodes=1)      # methane: one atom, no bonds
    single.ndata["h"] = torch.zeros(1, 117)
    two = pb.batch([single, single])
    assert two.number_of_nodes() == 2 and two.number_of_edges() == 0 and two.batch_size == 2
    c = two.csr()
    assert c.indptr.tolist() == [0, 0, 0] and c.has_zero_in_degree
    assert c.in_norm.tolist() == [1.0, 1.0]
    with pytest.raises(ValueError):
        BatchedGraph(torch.tensor([0]), torch.tensor([3]), 2)
    with pytest.raises(ValueError):
        two.ndata["h"] = torch.zeros(3, 4)


def test_dgl_style_construction_and_local_scope():
    g = pb.dgl.DGLGraph()
    g.add_nodes(3)
    g.ndata["h"] = torch.arange(6.0).reshape(3, 2)
    g.add_edges([0, 1], [1, 2], data={"type": torch.tensor([1.0, 2.0])})
    g.add_edges([1, 2], [0, 1], data={"type": torch.tensor([1.0, 2.0])})
    assert g.number_of_edges() == 4 and g.edata["type"].tolist() == [1.0, 2.0, 1.0, 2.0]
    with g.local_scope():
        g.ndata["h"] = torch.zeros(3, 2)
        g.ndata["tmp"] = torch.ones(3, 1)
    assert "tmp" not in g.ndata and g.ndata["h"][2, 1] == 5.0


def test_synthetic_shape_statistics():
    g, y = synthetic.make_batch(2000, seed=2666)
    assert y.shape == (2000, 1)
    assert 24.0 < g.number_of_nodes() / 2000 < 26.0
    assert 51.0 < g.number_of_edges() / 2000 < 57.0
    g2, y2 = synthetic.make_batch(2000, seed=2666)
    assert torch.equal(g.edges()[0], g2.edges()[0]) and torch.equal(y, y2)


def test_select_equals_unbatch_then_batch():
    """BatchedGraph.select == dgl.batch([dgl.unbatch(g)[i] for i in idxs]) bit for bit
    (pinot/active/experiment.py:83-104), checked against the oracle's unbatch / batch."""
    g, _ = synthetic.make_batch(40, seed=8)
    ref = dgl_ref.from_arrays(*g.edges(), g.number_of_nodes(), g.batch_num_nodes().tolist(),
                              g.batch_num_edges().tolist(), ndata={"h": g.ndata["h"]})
    ref_parts = dgl_ref.unbatch(ref)
    # (a `range` of consecutive graphs takes the slicing fast path, the lists the general gather)
    for idxs in ([5], [39, 0, 17, 17, 3], list(range(40)), list(range(10, 20)), range(10, 20), range(0, 40), range(39, 40),
                 range(0, 1)):
        got = g.select(idxs)
        want = dgl_ref.batch([ref_parts[i] for i in idxs])
        assert got.edges()[0].tolist() == want.src and got.edges()[1].tolist() == want.dst
        assert got.batch_num_nodes().tolist() == want.batch_nodes
        assert got.batch_num_edges().tolist() == want.batch_edges
        assert torch.equal(got.ndata["h"], want.ndata["h"])
    with pytest.raises(IndexError):
        g.select([40])
    from pinot_b200 import active
    gs, ys = active.slice_fn_tuple((g, torch.arange(40.0)[:, None]), [3, 1])
    assert gs.batch_size == 2 and ys.flatten().tolist() == [3.0, 1.0]
    with pytest.raises(RuntimeError, match="more than one"):
        active.slice_fn_graph(g.select([0]), [0])
