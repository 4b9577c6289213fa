"""Batched molecular graph: the stand-in for ``dgl.DGLGraph`` on the hot path.

Follows the layout produced by ``pinot.graph.from_rdkit_mol`` (pinot/graph.py:192-244: node
features ``ndata["h"]`` (n, 117), edges = all bonds begin->end followed by all bonds end->begin) and
the batching of ``dgl.batch`` as used by ``pinot.data.utils.batch`` (pinot/data/utils.py:235-282):
node ids are offset by the cumulative node count, per-graph edge blocks are concatenated in order,
and per-graph node / edge counts are retained.  Index tensors are int64 like DGL's; the int32 CSR
views consumed by the kernels are derived once per graph and cached.
"""
import contextlib

import os

import torch


_KERNEL_BUILD_MAX_NODES = 4096 * 16384   # limit of pinot_b200_csr_build's two-level scan


class GraphCSR:
    """Device-resident kernel view of a batched graph (built lazily, cached on the graph)."""

    def __init__(self, src, dst, num_nodes, batch_num_nodes):
        dev = src.device
        n = int(num_nodes)
        if src.numel() and (n >= 2 ** 31 or src.numel() >= 2 ** 31):
            raise ValueError("pinot_b200: graphs are limited to 2^31 nodes / edges (int32 CSR)")
        src = src.long()
        dst = dst.long()
        self.num_nodes = n
        self._has_zero_in_degree = None
        if dev.type == "cuda" and os.environ.get("PINOT_B200_CSR", "kernel") != "torch" and n <= _KERNEL_BUILD_MAX_NODES:
            self._build_with_kernel(src.contiguous(), dst.contiguous(), n, batch_num_nodes)
        else:
            self._build_with_torch(src, dst, n, batch_num_nodes)

    def _build_with_kernel(self, src, dst, n, batch_num_nodes):
        """One C-ABI call (pinot_b200_csr_build): counting sort with DGL's edge order kept inside rows,
        degree vectors, graph offsets; asynchronous on the current stream, no host synchronisation."""
        from . import _lib
        dev = src.device
        bnn = torch.as_tensor(batch_num_nodes, dtype=torch.long)
        self.batch_size = int(bnn.numel())
        if not bnn.is_cuda:
            bnn = bnn.pin_memory().to(dev, non_blocking=True)   # a pageable copy would block the host
        ne = int(src.numel())
        i32 = dict(dtype=torch.int32, device=dev)
        f32 = dict(dtype=torch.float32, device=dev)
        self.indptr, self.indices = torch.empty(n + 1, **i32), torch.empty(ne, **i32)
        self.indptr_t, self.indices_t = torch.empty(n + 1, **i32), torch.empty(ne, **i32)
        self.in_norm, self.out_norm, self.inv_in_deg = (torch.empty(n, **f32) for _ in range(3))
        self._in_deg = torch.empty(n, dtype=torch.long, device=dev)
        self.seg_offsets = torch.empty(self.batch_size + 1, **i32)
        self.node_seg = torch.empty(n, **i32)
        ws_bytes = int(_lib.load().pinot_b200_csr_build_workspace_bytes(n, ne))
        ws = torch.empty(ws_bytes // 4, **i32)
        _lib.call("csr_build", _lib.ptr(src), _lib.ptr(dst), ne, n, _lib.ptr(bnn), self.batch_size,
                  _lib.ptr(self.indptr), _lib.ptr(self.indices), _lib.ptr(self.indptr_t), _lib.ptr(self.indices_t),
                  _lib.ptr(self.in_norm), _lib.ptr(self.out_norm), _lib.ptr(self.inv_in_deg), _lib.ptr(self._in_deg),
                  _lib.ptr(self.seg_offsets), _lib.ptr(self.node_seg), _lib.ptr(ws), ws_bytes)

    def _build_with_torch(self, src, dst, n, batch_num_nodes):
        """The same index in torch ops (CPU graphs, and the cross-check of the kernel path)."""
        dev = src.device
        # degree counts by scatter-add: torch.bincount reads its maximum back to the host, which would
        # stall an asynchronous upload pipeline on every graph
        ones = torch.ones_like(dst)
        in_deg = torch.zeros(n, dtype=torch.long, device=dev).scatter_add_(0, dst, ones)
        out_deg = torch.zeros(n, dtype=torch.long, device=dev).scatter_add_(0, src, ones)
        # CSR by destination (forward gather) and by source (transposed graph, backward gather);
        # stable sorts keep DGL's edge order inside every row.
        order = torch.sort(dst, stable=True).indices
        self.indices = src[order].to(torch.int32).contiguous()
        self.indptr = torch.zeros(n + 1, dtype=torch.int32, device=dev)
        self.indptr[1:] = torch.cumsum(in_deg, 0).to(torch.int32)
        order_t = torch.sort(src, stable=True).indices
        self.indices_t = dst[order_t].to(torch.int32).contiguous()
        self.indptr_t = torch.zeros(n + 1, dtype=torch.int32, device=dev)
        self.indptr_t[1:] = torch.cumsum(out_deg, 0).to(torch.int32)
        # DGL GraphConv(norm="both"): degrees clamped to >= 1, then ^-1/2
        self.in_norm = in_deg.float().clamp(min=1).pow(-0.5).contiguous()
        self.out_norm = out_deg.float().clamp(min=1).pow(-0.5).contiguous()
        # SAGEConv("mean"): mean over in-neighbours, 0 for isolated nodes
        self.inv_in_deg = (1.0 / in_deg.float().clamp(min=1)).contiguous()
        bnn = torch.as_tensor(batch_num_nodes, dtype=torch.long)
        if dev.type == "cuda" and not bnn.is_cuda:
            # pinned + non_blocking: a pageable copy would block the host until the stream drains
            bnn = bnn.pin_memory().to(dev, non_blocking=True)
        else:
            bnn = bnn.to(dev)
        self.batch_size = int(bnn.numel())
        self.seg_offsets = torch.zeros(self.batch_size + 1, dtype=torch.int32, device=dev)
        self.seg_offsets[1:] = torch.cumsum(bnn, 0).to(torch.int32)
        self.node_seg = torch.repeat_interleave(
            torch.arange(self.batch_size, device=dev, dtype=torch.int32), bnn, output_size=n
        ).contiguous()
        self._in_deg = in_deg

    @property
    def has_zero_in_degree(self):
        """Evaluated on demand (one device->host read): only GraphConv(allow_zero_in_degree=False) asks."""
        if self._has_zero_in_degree is None:
            self._has_zero_in_degree = bool((self._in_deg == 0).any().item()) if self.num_nodes else False
        return self._has_zero_in_degree

    _TENSORS = ("indices", "indptr", "indices_t", "indptr_t", "in_norm", "out_norm", "inv_in_deg", "seg_offsets",
                "node_seg", "_in_deg")

    def copy_from_(self, other):
        """Overwrite this view's device arrays with another graph's (same node / edge / graph counts):
        the addresses stay, which is what a captured CUDA graph of the train step needs."""
        if (self.num_nodes, self.batch_size, self.indices.numel()) != (other.num_nodes, other.batch_size,
                                                                       other.indices.numel()):
            raise ValueError("copy_from_: graphs differ in size (%d nodes / %d edges / %d graphs vs %d / %d / %d)" % (
                self.num_nodes, self.indices.numel(), self.batch_size, other.num_nodes, other.indices.numel(),
                other.batch_size))
        for name in self._TENSORS:
            getattr(self, name).copy_(getattr(other, name), non_blocking=True)
        self._has_zero_in_degree = other._has_zero_in_degree
        return self

    def scales(self, norm):
        if norm == "both":
            return self.out_norm, self.in_norm
        if norm == "sym_in":     # SGConv / TAGConv: the in-degree normalises both sides
            return self.in_norm, self.in_norm
        if norm == "mean":
            return None, self.inv_in_deg
        if norm == "none":
            return None, None
        raise ValueError("unknown norm %r" % (norm,))


class _Frame(dict):
    """``g.ndata`` / ``g.edata``: a dict of per-node (per-edge) tensors with a length check."""

    def __init__(self, owner, count_fn):
        super().__init__()
        self._count_fn = count_fn

    def __setitem__(self, key, value):
        n = self._count_fn()
        if torch.is_tensor(value) and value.shape[0] != n:
            raise ValueError("expected %d rows for feature %r, got %d" % (n, key, value.shape[0]))
        super().__setitem__(key, value)


class BatchedGraph:
    """A batch of graphs with DGL-compatible surface (``ndata``, ``local_scope``, ``to``,
    ``batch_size``, ``number_of_nodes`` ...) used by pinot's representation, data and active-learning
    code (pinot/representation/sequential.py:78-103, pinot/data/utils.py:260-276,
    pinot/active/experiment.py:99-104)."""

    def __init__(self, src=None, dst=None, num_nodes=0, batch_num_nodes=None, batch_num_edges=None, validate=True,
                 batch_num_nodes_device=None):
        #: optional int64 device copy of batch_num_nodes (an upload pipeline copies it alongside the edges,
        #: so that building the kernel views needs no host->device transfer of its own)
        self._bnn_dev = batch_num_nodes_device
        self._src = torch.zeros(0, dtype=torch.long) if src is None else torch.as_tensor(src, dtype=torch.long)
        self._dst = torch.zeros(0, dtype=torch.long) if dst is None else torch.as_tensor(dst, dtype=torch.long).to(self._src.device)
        if self._src.shape != self._dst.shape:
            raise ValueError("src and dst must have the same length")
        self._num_nodes = int(num_nodes)
        # (validate=False with plain lists: an internal caller passing counts it took from another graph; converting
        #  65 536 of them one by one is milliseconds of host time per sub-batch of a pool)
        trusted = (not validate) and type(batch_num_nodes) is list and type(batch_num_edges) is list
        if trusted:
            self._bnn, self._bne = batch_num_nodes, batch_num_edges
        else:
            self._bnn = [self._num_nodes] if batch_num_nodes is None else [int(v) for v in batch_num_nodes]
            self._bne = [int(self._src.numel())] if batch_num_edges is None else [int(v) for v in batch_num_edges]
        if validate and (sum(self._bnn) != self._num_nodes or sum(self._bne) != self._src.numel()):
            raise ValueError("per-graph counts do not add up to the totals")
        if validate and self._src.numel():
            hi = int(max(self._src.max().item(), self._dst.max().item()))
            lo = int(min(self._src.min().item(), self._dst.min().item()))
            if lo < 0 or hi >= self._num_nodes:
                raise ValueError("edge endpoint out of range [0, %d)" % self._num_nodes)
        self.ndata = _Frame(self, lambda: self._num_nodes)
        self.edata = _Frame(self, lambda: int(self._src.numel()))
        self._csr = None
        self._off = None
        self._off_host = None

    # -- construction in the style of pinot/graph.py:208-239 -----------------------------------
    def add_nodes(self, n):
        if self._src.numel() or len(self._bnn) != 1:
            raise RuntimeError("add_nodes is only supported while building a single empty graph")
        self._num_nodes += int(n)
        self._bnn = [self._num_nodes]
        self._csr = None
        self._off = None
        self._off_host = None

    def add_edges(self, u, v, data=None):
        if len(self._bnn) != 1:
            raise RuntimeError("add_edges is only supported on a single (unbatched) graph")
        u = torch.as_tensor(u, dtype=torch.long, device=self._src.device).reshape(-1)
        v = torch.as_tensor(v, dtype=torch.long, device=self._src.device).reshape(-1)
        if u.numel() != v.numel():
            raise ValueError("add_edges: endpoint lists differ in length")
        if u.numel() and (int(min(u.min(), v.min())) < 0 or int(max(u.max(), v.max())) >= self._num_nodes):
            raise ValueError("add_edges: endpoint out of range")
        self._src = torch.cat([self._src, u])
        self._dst = torch.cat([self._dst, v])
        self._bne = [int(self._src.numel())]
        n_new = int(u.numel())
        n_old = int(self._src.numel()) - n_new
        for k in list(self.edata.keys()):
            old = self.edata[k]
            new = None if data is None else data.get(k)
            if new is None:
                new = torch.zeros((n_new,) + tuple(old.shape[1:]), dtype=old.dtype, device=old.device)
            dict.__setitem__(self.edata, k, torch.cat([old, torch.as_tensor(new, dtype=old.dtype)], dim=0))
        if data is not None:
            for k, new in data.items():
                if k not in self.edata:
                    new = torch.as_tensor(new)
                    pad = torch.zeros((n_old,) + tuple(new.shape[1:]), dtype=new.dtype)
                    dict.__setitem__(self.edata, k, torch.cat([pad, new], dim=0))
        self._csr = None
        self._off = None
        self._off_host = None

    # -- DGL surface ----------------------------------------------------------------------------
    @property
    def batch_size(self):
        return len(self._bnn)

    def number_of_nodes(self):
        return self._num_nodes

    num_nodes = number_of_nodes

    def number_of_edges(self):
        return int(self._src.numel())

    num_edges = number_of_edges

    def batch_num_nodes(self):
        return torch.tensor(self._bnn, dtype=torch.long)

    def batch_num_edges(self):
        return torch.tensor(self._bne, dtype=torch.long)

    def edges(self):
        return self._src, self._dst

    def all_edges(self):
        return self.edges()

    def in_degrees(self):
        return torch.bincount(self._dst, minlength=self._num_nodes)

    def out_degrees(self):
        return torch.bincount(self._src, minlength=self._num_nodes)

    @property
    def device(self):
        return self._src.device

    def to(self, device, **kwargs):
        g = BatchedGraph.__new__(BatchedGraph)
        g._src = self._src.to(device)
        g._dst = self._dst.to(device)
        g._num_nodes = self._num_nodes
        g._bnn = list(self._bnn)
        g._bne = list(self._bne)
        g.ndata = _Frame(g, lambda: g._num_nodes)
        g.edata = _Frame(g, lambda: int(g._src.numel()))
        for k, v in self.ndata.items():
            dict.__setitem__(g.ndata, k, v.to(device) if torch.is_tensor(v) else v)
        for k, v in self.edata.items():
            dict.__setitem__(g.edata, k, v.to(device) if torch.is_tensor(v) else v)
        g._csr = self._csr if (self._csr is not None and torch.device(device) == self._src.device) else None
        g._off = None
        g._off_host = None
        return g

    def copy_from_(self, other):
        """In-place copy of another batched graph with the same node / edge / graph counts (structure,
        kernel views and node / edge data); every device address of this graph stays valid."""
        if self._src.numel() != other._src.numel() or self._num_nodes != other._num_nodes or \
                len(self._bnn) != len(other._bnn):
            raise ValueError("copy_from_: graphs differ in size")
        self._src.copy_(other._src, non_blocking=True)
        self._dst.copy_(other._dst, non_blocking=True)
        self._bnn, self._bne = list(other._bnn), list(other._bne)
        self._off = None
        self._off_host = None
        self.csr().copy_from_(other.csr())
        for frame, oframe in ((self.ndata, other.ndata), (self.edata, other.edata)):
            for k, v in oframe.items():
                if torch.is_tensor(v) and k in frame and frame[k].shape == v.shape:
                    frame[k].copy_(v, non_blocking=True)
                else:
                    dict.__setitem__(frame, k, v.to(self._src.device) if torch.is_tensor(v) else v)
        return self

    @contextlib.contextmanager
    def local_scope(self):
        saved_n = dict(self.ndata)
        saved_e = dict(self.edata)
        try:
            yield self
        finally:
            dict.clear(self.ndata)
            dict.update(self.ndata, saved_n)
            dict.clear(self.edata)
            dict.update(self.edata, saved_e)

    def local_var(self):
        g = self.to(self.device)
        g._csr = self._csr
        return g

    def __repr__(self):
        return "BatchedGraph(batch_size=%d, num_nodes=%d, num_edges=%d, device=%s)" % (
            self.batch_size, self._num_nodes, self.number_of_edges(), self.device)

    # -- selection ------------------------------------------------------------------------------
    def _offsets(self):
        """(node offsets, edge offsets) per graph as int64 tensors on the graph's device (cached)."""
        if getattr(self, "_off", None) is None or self._off[0].device != self._src.device:
            dev = self._src.device
            bnn = torch.tensor(self._bnn, dtype=torch.long)
            bne = torch.tensor(self._bne, dtype=torch.long)
            noff = torch.zeros(len(self._bnn) + 1, dtype=torch.long)
            eoff = torch.zeros(len(self._bne) + 1, dtype=torch.long)
            noff[1:] = torch.cumsum(bnn, 0)
            eoff[1:] = torch.cumsum(bne, 0)
            self._off = (noff.to(dev), eoff.to(dev))
        return self._off

    def select(self, idxs):
        """``dgl.batch([dgl.unbatch(self)[i] for i in idxs])`` without materialising the individual
        graphs (pinot/active/experiment.py:83-104 re-batches like this every acquisition round):
        index arithmetic on the device, node / edge data moved by the row-gather kernel."""
        dev = self._src.device
        nb = len(self._bnn)
        if isinstance(idxs, range) and idxs.step == 1 and 0 <= idxs.start < idxs.stop <= nb:
            # A run of consecutive graphs (how a pool is scored in chunks): their nodes and edges are consecutive too,
            # so the sub-batch is slices -- node / edge data as views, no gather -- and one offset subtraction.
            # (The general path below spends ~16 ms of host time on 65 536 graphs building its index lists.)
            lo, hi = idxs.start, idxs.stop
            if getattr(self, "_off_host", None) is None or len(self._off_host[0]) != nb + 1:
                from itertools import accumulate
                self._off_host = ([0] + list(accumulate(self._bnn)), [0] + list(accumulate(self._bne)))
            n0, n1 = self._off_host[0][lo], self._off_host[0][hi]
            e0, e1 = self._off_host[1][lo], self._off_host[1][hi]
            out = BatchedGraph(self._src[e0:e1] - n0, self._dst[e0:e1] - n0, n1 - n0, self._bnn[lo:hi], self._bne[lo:hi],
                               validate=False)
            for name, frame, a, b in (("ndata", self.ndata, n0, n1), ("edata", self.edata, e0, e1)):
                for k, v in frame.items():
                    dict.__setitem__(getattr(out, name), k, v[a:b])
            return out
        idx_host = [int(i) for i in (idxs.tolist() if torch.is_tensor(idxs) else idxs)]
        idx_host = [i + nb if i < 0 else i for i in idx_host]
        if any(i < 0 or i >= nb for i in idx_host):
            raise IndexError("graph index out of range")
        if not idx_host:
            raise ValueError("select() needs at least one graph")
        idx = torch.tensor(idx_host, dtype=torch.long, device=dev)
        noff, eoff = self._offsets()
        cn = noff[idx + 1] - noff[idx]                      # nodes per selected graph
        ce = eoff[idx + 1] - eoff[idx]
        new_noff = torch.zeros(len(idx_host) + 1, dtype=torch.long, device=dev)
        new_eoff = torch.zeros_like(new_noff)
        new_noff[1:] = torch.cumsum(cn, 0)
        new_eoff[1:] = torch.cumsum(ce, 0)
        bnn = [self._bnn[i] for i in idx_host]
        bne = [self._bne[i] for i in idx_host]
        n_new, e_new = sum(bnn), sum(bne)
        # source position of every new node / edge
        which_n = torch.repeat_interleave(torch.arange(len(idx_host), device=dev), cn, output_size=n_new)
        node_src = torch.arange(n_new, device=dev) - new_noff[which_n] + noff[idx][which_n]
        which_e = torch.repeat_interleave(torch.arange(len(idx_host), device=dev), ce, output_size=e_new)
        edge_src = torch.arange(e_new, device=dev) - new_eoff[which_e] + eoff[idx][which_e]
        shift = (new_noff[:-1] - noff[idx])[which_e]        # old node id -> new node id, per edge
        out = BatchedGraph(self._src[edge_src] + shift, self._dst[edge_src] + shift, n_new, bnn, bne, validate=False)
        for name, frame, index in (("ndata", self.ndata, node_src), ("edata", self.edata, edge_src)):
            for k, v in frame.items():
                dict.__setitem__(getattr(out, name), k, _gather_rows(v, index))
        return out

    # -- kernel view ----------------------------------------------------------------------------
    def csr(self):
        if self._csr is None:
            bnn_dev = getattr(self, "_bnn_dev", None)
            use_dev = bnn_dev is not None and bnn_dev.device == self._src.device and bnn_dev.numel() == len(self._bnn)
            self._csr = GraphCSR(self._src, self._dst, self._num_nodes, bnn_dev if use_dev else self._bnn)
        return self._csr


def _gather_rows(v, index):
    """v[index] along dim 0; fp32 CUDA matrices go through the library's row-gather kernel."""
    if v.is_cuda and v.dtype == torch.float32 and v.dim() == 2 and v.is_contiguous():
        from ._lib import call, ptr
        out = torch.empty((index.numel(), v.shape[1]), device=v.device, dtype=torch.float32)
        call("gather_rows", ptr(v), ptr(index), ptr(out), index.numel(), v.shape[1])
        return out
    return v[index]


def batch(graphs):
    """``dgl.batch``: concatenate graphs; node ids of graph k are offset by the nodes before it."""
    graphs = list(graphs)
    if not graphs:
        raise ValueError("batch() needs at least one graph")
    srcs, dsts, bnn, bne = [], [], [], []
    offset = 0
    for g in graphs:
        srcs.append(g._src + offset)
        dsts.append(g._dst + offset)
        bnn.extend(g._bnn)
        bne.extend(g._bne)
        offset += g._num_nodes
    out = BatchedGraph(torch.cat(srcs), torch.cat(dsts), offset, bnn, bne)
    for frame_name in ("ndata", "edata"):
        keys = set(getattr(graphs[0], frame_name).keys())
        for g in graphs[1:]:
            keys &= set(getattr(g, frame_name).keys())
        for k in sorted(keys):
            dict.__setitem__(getattr(out, frame_name), k,
                             torch.cat([getattr(g, frame_name)[k] for g in graphs], dim=0))
    return out


def unbatch(g):
    """``dgl.unbatch``: the inverse of :func:`batch`."""
    out = []
    n0 = e0 = 0
    for n, e in zip(g._bnn, g._bne):
        sub = BatchedGraph(g._src[e0:e0 + e] - n0, g._dst[e0:e0 + e] - n0, n)
        for k, v in g.ndata.items():
            dict.__setitem__(sub.ndata, k, v[n0:n0 + n])
        for k, v in g.edata.items():
            dict.__setitem__(sub.edata, k, v[e0:e0 + e])
        out.append(sub)
        n0 += n
        e0 += e
    return out


def sum_nodes(g, feat):
    """``dgl.sum_nodes(g, feat)``: per-graph sum of a node feature -> (batch_size, F)."""
    from . import ops
    return ops.segment_readout(g.ndata[feat], g.csr(), mean=False)


def mean_nodes(g, feat):
    """``dgl.mean_nodes(g, feat)``."""
    from . import ops
    return ops.segment_readout(g.ndata[feat], g.csr(), mean=True)
