"""pinot_b200.data.PackedDataset / PackedBatch: the compact host form reproduces the dense
features bit for bit, refuses what it cannot reproduce, and the container behaves like the reference's
``pinot.data.datasets.Dataset`` (pinot/data/datasets.py:17-113) where a training loop can tell."""
import random

import pytest
import torch

import pinot_b200 as pb
from pinot_b200 import synthetic
from pinot_b200.data import PackedBatch, PackedDataset, compact_features, expand_features_host

DEV = "cuda"


def _pairs(n_entries, mols_per_entry, seed=0):
    out = []
    for k in range(n_entries):
        g, y = synthetic.make_batch(mols_per_entry, seed=seed + k)
        out.append((g, y))
    return out


def _reference_like(g):
    """The synthetic generator writes an atomic MASS into the isotope column; the reference's featurisers write
    the isotope number, an integer (pinot/graph.py:57, :117).  Integers and flags only, as the reference emits."""
    h = g.ndata["h"].clone()
    h[:, 105] = 0.0
    g.ndata["h"] = h
    return g


def test_compact_form_is_bit_exact_and_thirteen_times_smaller():
    g, y = synthetic.make_batch(64, seed=3)
    g = _reference_like(g)
    h = g.ndata["h"]
    z, fp = compact_features(h)
    assert z.dtype == torch.uint8 and fp.dtype == torch.float16 and fp.shape == (h.shape[0], 17)
    assert torch.equal(expand_features_host(z, fp), h)
    e = PackedBatch.from_graph(g, y)
    assert e.compact and len(e) == 64
    assert torch.equal(e.features(), h)
    dense_bytes = h.numel() * 4
    compact_bytes = e.z.numel() + 2 * e.fp.numel()
    assert dense_bytes / compact_bytes > 13
    g2 = e.graph()
    assert torch.equal(g2.edges()[0], g.edges()[0]) and torch.equal(g2.edges()[1], g.edges()[1])
    assert g2.batch_num_nodes().tolist() == g.batch_num_nodes().tolist()
    assert torch.equal(g2.ndata["h"], h)


@pytest.mark.parametrize("spoil", ["fraction", "large", "nan"])
def test_columns_fp16_cannot_hold_travel_as_fp32(spoil):
    g, y = synthetic.make_batch(8, seed=5)
    h = _reference_like(g).ndata["h"].clone()
    if spoil == "fraction":
        h[3, 105] = 12.0107            # an atomic mass: not an fp16 number (the synthetic generator's own column 105)
    elif spoil == "large":
        h[0, 101] = 70000.0            # beyond fp16
    else:
        h[0, 110] = float("nan")
    z, fp = compact_features(h)
    assert fp.dtype == torch.float32
    back = expand_features_host(z, fp)
    assert torch.equal(torch.nan_to_num(back, nan=-7.0), torch.nan_to_num(h, nan=-7.0))
    g.ndata["h"] = h
    e = PackedBatch.from_graph(g, y)
    assert e.compact and h.numel() * 4 / (e.z.numel() + 4 * e.fp.numel()) > 6.5
    # the unmodified synthetic batches are of this kind
    e2 = PackedBatch.from_graph(*synthetic.make_batch(8, seed=6))
    assert e2.compact and e2.fp.dtype == torch.float32


@pytest.mark.parametrize("spoil", ["two_hot", "no_hot", "half_hot"])
def test_features_without_an_exact_one_hot_stay_dense(spoil):
    g, y = synthetic.make_batch(8, seed=5)
    h = g.ndata["h"].clone()
    if spoil == "two_hot":
        h[2, :100] = 0.0
        h[2, 7] = 1.0
        h[2, 8] = 1.0
    elif spoil == "no_hot":
        h[1, :100] = 0.0
    else:
        h[1, :100] = 0.0
        h[1, 6] = 0.5
        h[1, 8] = 0.5
    assert compact_features(h) is None
    g.ndata["h"] = h
    e = PackedBatch.from_graph(g, y)
    assert not e.compact and torch.equal(e.features(), h)
    assert e.transfer_bytes() > h.numel() * 4


def test_container_semantics_follow_the_reference_dataset():
    pairs = _pairs(10, 3)
    ds = PackedDataset.from_pairs(pairs)
    assert len(ds) == 30                                   # molecules, not entries (datasets.py:25-31)
    assert isinstance(ds[2], PackedBatch) and isinstance(ds[2:5], PackedDataset) and len(ds[2:5].ds) == 3
    assert [len(e) for e in ds] == [3] * 10
    with pytest.raises(RuntimeError):
        PackedDataset()[0]
    assert len(PackedDataset()) == 0
    assert len((ds[:2] + ds[5:]).ds) == 7
    # shuffle: Python's random, reseeded, in place (datasets.py:64-69)
    order = [id(e) for e in ds.ds]
    expect = list(order)
    random.seed(11)
    random.shuffle(expect)
    assert [id(e) for e in ds.shuffle(seed=11).ds] == expect
    # split: sizes from len(ds) = MOLECULES, applied to ENTRIES, truncated (utils.split through datasets.py:51-62)
    ds = PackedDataset.from_pairs(_pairs(10, 1))
    a, b = ds.split([4, 1], seed=0)
    assert (len(a.ds), len(b.ds)) == (8, 2)
    ds3 = PackedDataset.from_pairs(_pairs(10, 1))
    parts = ds3.split([1, 1, 1], seed=0)
    assert [len(p.ds) for p in parts] == [3, 3, 3]          # one entry is left out, as in the reference


def test_batch_matches_utils_batch_on_the_same_entries():
    pairs = _pairs(12, 1, seed=20)
    ds = PackedDataset.from_pairs(pairs)
    out = ds.batch(5)
    want = pb.data.utils.batch(list(pairs), 5)
    assert len(out.ds) == len(want) == 2                    # 12 // 5, remainder dropped
    for e, (g, y) in zip(out.ds, want):
        assert torch.equal(e.features(), g.ndata["h"]) and torch.equal(e.y, y)
        assert torch.equal(e.src, g.edges()[0]) and e.batch_num_nodes == g.batch_num_nodes().tolist()
    assert len(ds.batch(5, partial_batch=True).ds) == 3
    # on the CPU, `to` hands back host graphs (no device, no kernels)
    (g0, y0), = PackedDataset(out.ds[:1]).to("cpu")
    assert torch.equal(g0.ndata["h"], want[0][0].ndata["h"]) and torch.equal(y0, want[0][1])


@pytest.mark.gpu
def test_device_expansion_equals_the_dense_matrix():
    g, y = synthetic.make_batch(300, seed=9)
    for gg in (g, _reference_like(synthetic.make_batch(300, seed=9)[0])):      # fp32 columns, fp16 columns
        h = gg.ndata["h"]
        z, fp = compact_features(h)
        got = pb.ops.expand_atom_features(z.to(DEV), fp.to(DEV))
        assert got.dtype == torch.float32 and torch.equal(got.cpu(), h)
    # other widths, empty input
    z2 = torch.randint(0, 7, (50,), dtype=torch.uint8)
    fp2 = torch.randn(50, 3).to(torch.float16)
    got2 = pb.ops.expand_atom_features(z2.to(DEV), fp2.to(DEV), n_onehot=7).cpu()
    assert torch.equal(got2, expand_features_host(z2, fp2, 7))
    assert pb.ops.expand_atom_features(z2[:0].to(DEV), fp2[:0].to(DEV), n_onehot=7).shape == (0, 10)


@pytest.mark.gpu
def test_packed_entry_on_the_device_equals_the_dense_upload():
    g, y = synthetic.make_batch(200, seed=13)
    e = PackedBatch.from_graph(g, y)
    gd, yd = e.to(DEV)
    ref = g.to(DEV)
    assert torch.equal(gd.ndata["h"], ref.ndata["h"]) and torch.equal(yd.cpu(), y)
    for name in ("indptr", "indices", "indptr_t", "indices_t"):
        assert torch.equal(getattr(gd.csr(), name), getattr(ref.csr(), name))
    torch.manual_seed(0)
    net = pb.Net(pb.representation.Sequential(pb.representation.dgl_legacy.gn("GraphConv"), [32, "tanh", 32, "tanh"]),
                 output_regressor_class=pb.regressors.ExactGaussianProcessRegressor).to(DEV)
    # (the packed upload lays the feature rows out with a pitch of 120 floats, so its first node update multiplies a
    #  zero-padded matrix: the same sums, possibly by another kernel -- equal to rounding, not bit for bit)
    assert gd.ndata["h"].stride() == (120, 1)
    assert torch.allclose(net.representation(gd), net.representation(ref), rtol=1e-5, atol=1e-6)
    # the captured upload takes the compact tensors and yields the same graph
    src_h, dst_h, feats, y_h = e.host_tensors()
    up = pb.CapturedBatchUpload(src_h, dst_h, (feats[0], feats[1], e.n_onehot), y_h, e.num_nodes, e.batch_num_nodes,
                                e.batch_num_edges, torch.cuda.Stream())
    g2, y2, ev = up.replay()
    ev.synchronize()
    assert torch.equal(g2.ndata["h"], ref.ndata["h"]) and torch.equal(y2.cpu(), y)
    assert up.h2d_bytes < 0.25 * (g.ndata["h"].numel() * 4)
    up.release()


@pytest.mark.gpu
def test_padded_pitch_features_give_the_same_first_layer():
    """expand_atom_features(pad_pitch=True): same values, rows 16-byte aligned; node_update on the padded view (the
    TMA-fed kernel at this size) against the dense layout (register-staged kernel), forward and weight gradients."""
    torch.manual_seed(3)
    n = 20000
    z = torch.randint(0, 100, (n,), dtype=torch.uint8, device=DEV)
    fp = torch.randint(0, 5, (n, 17), device=DEV).to(torch.float16)
    dense = pb.ops.expand_atom_features(z, fp)
    padded = pb.ops.expand_atom_features(z, fp, pad_pitch=True)
    assert dense.is_contiguous() and padded.shape == dense.shape and padded.stride() == (120, 1)
    assert torch.equal(dense, padded) and float(padded._pb_pitch_base[:, 117:].abs().max()) == 0.0
    outs = []
    for h in (dense, padded):
        w = (torch.randn(128, 117, device=DEV, generator=torch.Generator(DEV).manual_seed(5)) * 0.1).requires_grad_()
        b = torch.zeros(128, device=DEV).add_(0.01).requires_grad_()
        y = pb.ops.node_update(h, w, b, w_is_out_in=True, act="tanh")
        (y * torch.linspace(-1, 1, 128, device=DEV)).sum().backward()
        outs.append((y.detach(), w.grad, b.grad))
    for got, want in zip(outs[1], outs[0]):
        assert got.shape == want.shape
        assert float((got - want).norm() / want.norm()) < 2e-6, float((got - want).norm() / want.norm())
    # (in, out) weights (GraphConv's layout) and a view that is NOT the one the expansion handed out: dense path
    w2 = torch.randn(117, 32, device=DEV) * 0.1
    y_a = pb.ops.node_update(padded, w2, None, w_is_out_in=False)
    y_b = pb.ops.node_update(dense, w2, None, w_is_out_in=False)
    assert float((y_a - y_b).norm() / y_b.norm()) < 2e-6
    y_c = pb.ops.node_update(padded[:1000], w2, None, w_is_out_in=False)       # a slice carries no marker
    assert float((y_c - y_b[:1000]).norm() / y_b[:1000].norm()) < 2e-6


def test_save_and_load_round_trip(tmp_path):
    ds = PackedDataset.from_pairs(_pairs(4, 5, seed=30))
    path = str(tmp_path / "pool.pt")
    ds.save(path)
    back = PackedDataset().load(path)
    assert len(back) == len(ds) == 20 and len(back.ds) == 4
    for a, b in zip(ds.ds, back.ds):
        assert a.compact == b.compact and torch.equal(a.features(), b.features()) and torch.equal(a.y, b.y)
        assert torch.equal(a.src, b.src) and torch.equal(a.dst, b.dst) and a.batch_num_nodes == b.batch_num_nodes
    sub = PackedDataset().load(path, indices=[2, 0])
    assert len(sub.ds) == 2 and torch.equal(sub.ds[0].y, ds.ds[2].y)
    with pytest.raises(ValueError):
        torch.save({"format": "something else"}, path)
        PackedDataset().load(path)
