"""A ``Dataset``-like container that keeps molecules in a transfer-ready form.

The reference's ``pinot.data.datasets.Dataset`` (pinot/data/datasets.py:17-113) is a list of ``(DGLGraph, y)``
entries; ``.to(device)`` moves every graph with its dense ``(n_atoms, 117)`` fp32 feature matrix
(pinot/graph.py:163-173: a one-hot of the atomic number followed by 17 small integers and flags).  Here an entry is
a :class:`PackedBatch`: the batched edge list, the per-graph counts and the features in COMPACT form -- the atomic
number as one byte and the 17 fingerprint columns as fp16 (fp32 if a value does not survive fp16) -- in pinned
host memory.  ``to(device)`` ships 35 (or 69) bytes per atom instead of 468 and rebuilds the dense matrix on the
device (``pinot_b200_expand_atom_features``), then builds the graph index with one ``pinot_b200_csr_build`` call.

The compact form is only used when it reproduces the dense matrix BIT FOR BIT (checked when an entry is packed:
exactly one 1.0 among the first ``n_onehot`` columns; the other columns go through fp16 only if that round trip
changes nothing); otherwise the entry keeps its dense features and is shipped as the reference would ship it.

``PackedDataset`` mirrors the reference class where its behaviour decides what a training loop sees:
``len`` counts molecules (datasets.py:25-31), integer / slice indexing (:33-42), iteration (:44-49), ``shuffle``
reseeds Python's ``random`` and shuffles in place (:64-69), ``split`` = shuffle + ``utils.split`` (:51-62),
``batch`` (:71-74), ``save`` / ``load`` (:76-103, one file holding the packed arrays), ``+`` (:105-108), ``to``
(:110-113)."""
import random

import torch

from .. import graph as _graph
from . import utils as _utils

N_ONEHOT = 100      # pinot/graph.py:163
N_FP = 17           # pinot/graph.py:50-69 (12 integers / flags + 5 hybridisation flags)


def _pin(t):
    t = t.contiguous()
    return t.pin_memory() if torch.cuda.is_available() and not t.is_pinned() else t


def compact_features(h, n_onehot=N_ONEHOT):
    """``(z, fp)`` -- uint8 atomic numbers and the remaining columns, as fp16 when every value survives the round trip
    (integers and flags, as the reference's featurisers emit) and as fp32 otherwise -- if the first ``n_onehot``
    columns are an exact one-hot; else None."""
    if h.dim() != 2 or h.dtype != torch.float32 or h.shape[1] < n_onehot or n_onehot > 256:
        return None
    onehot, rest = h[:, :n_onehot], h[:, n_onehot:]
    z = onehot.argmax(dim=1)
    if not (bool(((onehot == 0) | (onehot == 1)).all()) and bool((onehot.sum(dim=1) == 1).all())):
        return None
    fp = rest.to(torch.float16)
    if not bool((fp.to(torch.float32) == rest).all()):     # a fraction, a value beyond fp16, a NaN: keep fp32
        fp = rest.clone()
    return z.to(torch.uint8), fp


def expand_features_host(z, fp, n_onehot=N_ONEHOT):
    """Host restatement of the device expansion: the dense matrix the reference would have built."""
    h = torch.zeros(z.shape[0], n_onehot + fp.shape[1], dtype=torch.float32)
    h[torch.arange(z.shape[0]), z.long()] = 1.0
    h[:, n_onehot:] = fp.to(torch.float32)
    return h            # (NaN-safe: assignment, no arithmetic)


class PackedBatch:
    """One batch of molecules on the host: edge list, counts, targets, features compact (``z``, ``fp``) or dense."""

    def __init__(self, src, dst, num_nodes, batch_num_nodes, batch_num_edges, y, z=None, fp=None, h=None, n_onehot=N_ONEHOT):
        assert (z is None) == (fp is None) and (h is None) != (z is None), "features: either (z, fp) or dense h"
        self.src, self.dst = _pin(src.to(torch.long)), _pin(dst.to(torch.long))
        self.num_nodes = int(num_nodes)
        self.batch_num_nodes = [int(v) for v in batch_num_nodes]
        self.batch_num_edges = [int(v) for v in batch_num_edges]
        self.y = _pin(y)
        self.z = None if z is None else _pin(z)
        self.fp = None if fp is None else _pin(fp)
        self.h = None if h is None else _pin(h)
        self.n_onehot = int(n_onehot)

    # ---- construction ------------------------------------------------------------------------------
    @classmethod
    def from_graph(cls, g, y, n_onehot=N_ONEHOT):
        """Pack a host ``BatchedGraph`` with ``ndata['h']`` and its targets."""
        h = g.ndata["h"].detach().to("cpu", torch.float32)
        packed = compact_features(h, n_onehot)
        kw = dict(h=h) if packed is None else dict(z=packed[0], fp=packed[1])
        return cls(g._src.cpu(), g._dst.cpu(), g.number_of_nodes(), g._bnn, g._bne, y.detach().cpu(), n_onehot=n_onehot, **kw)

    # ---- what a consumer of the reference's entries would ask for ------------------------------------
    @property
    def compact(self):
        return self.z is not None

    def __len__(self):
        return len(self.batch_num_nodes)

    def features(self):
        """The dense host feature matrix (what the reference keeps in ``g.ndata['h']``)."""
        return self.h if self.h is not None else expand_features_host(self.z, self.fp, self.n_onehot)

    def graph(self):
        """The host ``BatchedGraph`` this entry stands for."""
        g = _graph.BatchedGraph(self.src, self.dst, self.num_nodes, self.batch_num_nodes, self.batch_num_edges, validate=False)
        g.ndata["h"] = self.features()
        return g

    def host_tensors(self):
        """The tensors a transfer copies, in the order ``CapturedBatchUpload`` takes them."""
        feats = (self.z, self.fp) if self.compact else self.h
        return self.src, self.dst, feats, self.y

    def transfer_bytes(self):
        ts = [self.src, self.dst, self.y] + ([self.z, self.fp] if self.compact else [self.h])
        return sum(t.numel() * t.element_size() for t in ts) + 8 * len(self.batch_num_nodes)

    def to(self, device, non_blocking=True):
        """``(BatchedGraph, y)`` on ``device``: compact upload, device-side expansion, one index build."""
        from .. import ops
        device = torch.device(device)
        if device.type != "cuda":
            return self.graph(), self.y
        src, dst = self.src.to(device, non_blocking=non_blocking), self.dst.to(device, non_blocking=non_blocking)
        y = self.y.to(device, non_blocking=non_blocking)
        if self.compact:
            h = ops.expand_atom_features(self.z.to(device, non_blocking=non_blocking),
                                         self.fp.to(device, non_blocking=non_blocking), self.n_onehot, pad_pitch=True)
        else:
            h = self.h.to(device, non_blocking=non_blocking)
        g = _graph.BatchedGraph(src, dst, self.num_nodes, self.batch_num_nodes, self.batch_num_edges, validate=False)
        g.ndata["h"] = h
        g.csr()
        return g, y


class PackedDataset:
    """List of ``PackedBatch`` entries with the reference ``Dataset``'s list semantics."""

    def __init__(self, ds=None):
        self.ds = ds
        self.device = torch.device("cpu")

    @classmethod
    def from_pairs(cls, pairs, n_onehot=N_ONEHOT):
        """From the reference's representation: an iterable of ``(graph, y)`` entries."""
        return cls([PackedBatch.from_graph(g, y, n_onehot) for g, y in pairs])

    def __len__(self):
        return 0 if self.ds is None else sum(len(e) for e in self.ds)        # molecules, as datasets.py:25-31

    def __getitem__(self, idx):
        if self.ds is None:
            raise RuntimeError("Empty molecule dataset.")
        if isinstance(idx, int):
            return self.ds[idx]
        if isinstance(idx, slice):
            return self.__class__(ds=self.ds[idx])
        raise TypeError("PackedDataset indices are integers or slices")

    def __iter__(self):
        return iter(self.ds)

    def __add__(self, other):
        return self.__class__(self.ds + other.ds)

    def shuffle(self, seed=None):
        random.seed(seed)
        random.shuffle(self.ds)
        return self

    def split(self, partition, seed=None):
        """Shuffle, then ``utils.split``: sizes are computed from ``len(self)`` (MOLECULES, datasets.py:25-31) and
        truncated, but applied to the list of entries (datasets.py:51-62 with data/utils.py:219-230)."""
        self.shuffle(seed=seed)
        return tuple(_utils.split(self, partition))

    def batch(self, batch_size, seed=2666, shuffle=False, partial_batch=False):
        """Re-batch into entries of ``batch_size`` molecules with the semantics of ``utils.batch``."""
        pairs = [(e.graph(), e.y) for e in self.ds]
        out = _utils.batch(pairs, batch_size, seed=seed, shuffle=shuffle, partial_batch=partial_batch)
        n_onehot = self.ds[0].n_onehot if self.ds else N_ONEHOT
        return self.__class__([PackedBatch.from_graph(g, y, n_onehot) for g, y in out])

    def transfer_bytes(self):
        return sum(e.transfer_bytes() for e in self.ds)

    # ---- on-disk form (the reference's Dataset.save / load, datasets.py:76-103, go through dgl.save_graphs) ---------
    def save(self, path):
        """Write the entries as they are held -- compact features, edge lists, counts, targets -- to one file."""
        torch.save({"format": "pinot_b200.PackedDataset/1",
                    "entries": [{"src": e.src, "dst": e.dst, "num_nodes": e.num_nodes, "batch_num_nodes": e.batch_num_nodes,
                                 "batch_num_edges": e.batch_num_edges, "y": e.y, "z": e.z, "fp": e.fp, "h": e.h,
                                 "n_onehot": e.n_onehot} for e in (self.ds or [])]}, path)

    def load(self, path, indices=None):
        """Read a file written by :meth:`save` (``indices``: a subset of its entries, as in the reference)."""
        blob = torch.load(path, map_location="cpu", weights_only=False)
        if blob.get("format") != "pinot_b200.PackedDataset/1":
            raise ValueError("not a PackedDataset file: %r" % (path,))
        entries = blob["entries"]
        if indices is not None:
            entries = [entries[i] for i in indices]
        self.ds = [PackedBatch(d["src"], d["dst"], d["num_nodes"], d["batch_num_nodes"], d["batch_num_edges"], d["y"],
                               z=d["z"], fp=d["fp"], h=d["h"], n_onehot=d["n_onehot"]) for d in entries]
        return self

    def to(self, device):
        """Device-resident ``(graph, y)`` entries (a plain list: what a training loop iterates over)."""
        self.device = torch.device(device)
        return [e.to(self.device) for e in self.ds]
