"""pinot_b200_csr_build against the torch construction of the same index: indices and offsets bit-exact
(DGL's edge order inside every row), degree vectors to the last bit or ulp."""
import os

import pytest
import torch

from pinot_b200 import graph as pbg, synthetic

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _both(src, dst, n, bnn):
    src, dst = src.to(DEV), dst.to(DEV)
    k = pbg.GraphCSR(src, dst, n, bnn)
    t = pbg.GraphCSR.__new__(pbg.GraphCSR)
    t.num_nodes, t._has_zero_in_degree = n, None
    t._build_with_torch(src.long(), dst.long(), n, bnn)
    torch.cuda.synchronize()
    return k, t


def _check(k, t):
    for name in ("indptr", "indices", "indptr_t", "indices_t", "seg_offsets", "node_seg", "_in_deg"):
        a, b = getattr(k, name), getattr(t, name)
        assert a.dtype == b.dtype and a.shape == b.shape, name
        assert torch.equal(a, b), name
    assert torch.equal(k.inv_in_deg, t.inv_in_deg)
    for name in ("in_norm", "out_norm"):
        assert torch.allclose(getattr(k, name), getattr(t, name), rtol=2e-7, atol=0), name
    assert k.batch_size == t.batch_size and k.has_zero_in_degree == t.has_zero_in_degree


@pytest.mark.parametrize("n_mol,seed", [(1, 0), (37, 1), (4200, 2), (60000, 3)])
def test_synthetic_batches(n_mol, seed):
    g, _ = synthetic.make_batch(n_mol, seed=seed)
    s, d = g.edges()
    _check(*_both(s, d, g.number_of_nodes(), g.batch_num_nodes().tolist()))


def test_random_multigraph_with_isolated_nodes_empty_graphs_and_hubs():
    gen = torch.Generator().manual_seed(5)
    n = 9001
    src = torch.randint(0, n, (50000,), generator=gen)
    dst = torch.randint(0, n // 3, (50000,), generator=gen)          # two thirds of the nodes have no in-edges
    dst[:3000] = 7                                                   # a hub row of 3000+ entries
    src[40000:] = src[39999]                                         # repeated edges
    bnn = [0, 4000, 0, 0, 5000, 1, 0]                                # empty graphs in front, middle and end
    k, t = _both(src, dst, n, bnn)
    _check(k, t)
    assert k.has_zero_in_degree


def test_no_edges_and_no_nodes():
    e = torch.zeros(0, dtype=torch.long)
    _check(*_both(e, e, 5, [2, 3]))
    _check(*_both(e, e, 0, []))


def test_batched_graph_uses_the_kernel_and_the_layers_agree():
    import pinot_b200 as pb
    g, _ = synthetic.make_batch(300, seed=9)
    g = g.to(DEV)
    torch.manual_seed(0)
    conv = pb.dgl.nn.pytorch.conv.GraphConv(117, 32).to(DEV)
    out_kernel = conv(g, g.ndata["h"])
    os.environ["PINOT_B200_CSR"] = "torch"
    try:
        g2 = g.to(DEV)
        g2._csr = None
        out_torch = conv(g2, g2.ndata["h"])
    finally:
        del os.environ["PINOT_B200_CSR"]
    assert torch.allclose(out_kernel, out_torch, rtol=1e-6, atol=1e-6)
