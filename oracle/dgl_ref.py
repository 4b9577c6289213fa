"""Plain-torch CPU restatement of the DGL API surface pinot uses.  ORACLE -- test infrastructure only.

Restated from DGL's documented semantics (DGL 0.4.3-0.6.x; the source is not in /root/reference),
following pinot's call sites: pinot/graph.py:208-239 (DGLGraph, add_nodes, add_edges),
pinot/data/utils.py:260-276 (batch / unbatch), pinot/representation/sequential.py:78-103
(local_scope, sum_nodes), pinot/representation/dgl_legacy.py:40-50 (dgl.nn.pytorch.conv.<Layer>).

GraphConv(norm="both"): feat * out_deg.clamp(min=1)^-1/2 ; if in > out: (feat @ W) summed over
in-edges, else summed then @ W ; * in_deg.clamp(min=1)^-1/2 ; + bias.  weight (in, out),
xavier_uniform_; bias zeros.
SAGEConv("mean"): fc_self(feat) + fc_neigh(mean over in-edges of feat) (mean of nothing = 0); both
Linear(in, out, bias=True), xavier_uniform_ with gain("relu").
sum_nodes: per-graph sum in node order.  batch: node ids offset by the node counts of the preceding
graphs, per-graph edge lists concatenated in order.

Written independently of pinot_b200 (per-graph Python loops, index_add_) so that it can check it.
"""
import contextlib
import sys
import types

import torch


class DGLGraph:
    def __init__(self):
        self.n = 0
        self.src = []
        self.dst = []
        self.ndata = {}
        self.edata = {}
        self.batch_nodes = None  # list of per-graph node counts when batched
        self.batch_edges = None

    # construction ---------------------------------------------------------------------------
    def add_nodes(self, n):
        self.n += int(n)

    def add_edges(self, u, v, data=None):
        u = [int(a) for a in (u.tolist() if torch.is_tensor(u) else u)]
        v = [int(a) for a in (v.tolist() if torch.is_tensor(v) else v)]
        assert len(u) == len(v)
        self.src += u
        self.dst += v

    # queries --------------------------------------------------------------------------------
    def number_of_nodes(self):
        return self.n

    def number_of_edges(self):
        return len(self.src)

    @property
    def batch_size(self):
        return 1 if self.batch_nodes is None else len(self.batch_nodes)

    def node_counts(self):
        return [self.n] if self.batch_nodes is None else list(self.batch_nodes)

    def edge_counts(self):
        return [len(self.src)] if self.batch_edges is None else list(self.batch_edges)

    def in_degrees(self):
        return torch.bincount(torch.tensor(self.dst, dtype=torch.long), minlength=self.n)

    def out_degrees(self):
        return torch.bincount(torch.tensor(self.src, dtype=torch.long), minlength=self.n)

    def to(self, device, **kw):
        return self

    @contextlib.contextmanager
    def local_scope(self):
        saved = dict(self.ndata)
        try:
            yield self
        finally:
            self.ndata.clear()
            self.ndata.update(saved)

    def dense_adjacency(self, dtype=torch.float64):
        """A[i, j] = number of edges j -> i."""
        a = torch.zeros(self.n, self.n, dtype=dtype)
        for s, d in zip(self.src, self.dst):
            a[d, s] += 1
        return a


def from_arrays(src, dst, num_nodes, batch_num_nodes=None, batch_num_edges=None, ndata=None):
    g = DGLGraph()
    g.add_nodes(num_nodes)
    g.add_edges(src, dst)
    if batch_num_nodes is not None and len(batch_num_nodes) > 1:
        g.batch_nodes = [int(a) for a in batch_num_nodes]
        g.batch_edges = [int(a) for a in batch_num_edges]
    for k, v in (ndata or {}).items():
        g.ndata[k] = v
    return g


def batch(graphs):
    out = DGLGraph()
    out.batch_nodes, out.batch_edges = [], []
    for g in graphs:
        off = out.n
        out.n += g.n
        out.src += [s + off for s in g.src]
        out.dst += [d + off for d in g.dst]
        out.batch_nodes += g.node_counts()
        out.batch_edges += g.edge_counts()
    keys = set(graphs[0].ndata)
    for g in graphs[1:]:
        keys &= set(g.ndata)
    for k in keys:
        out.ndata[k] = torch.cat([g.ndata[k] for g in graphs], dim=0)
    return out


def unbatch(g):
    outs = []
    n0 = e0 = 0
    for n, e in zip(g.node_counts(), g.edge_counts()):
        s = DGLGraph()
        s.add_nodes(n)
        s.add_edges([a - n0 for a in g.src[e0:e0 + e]], [a - n0 for a in g.dst[e0:e0 + e]])
        for k, v in g.ndata.items():
            s.ndata[k] = v[n0:n0 + n]
        outs.append(s)
        n0 += n
        e0 += e
    return outs


def sum_nodes(g, feat):
    x = g.ndata[feat]
    rows = []
    n0 = 0
    for n in g.node_counts():
        rows.append(x[n0:n0 + n].sum(dim=0))
        n0 += n
    return torch.stack(rows, dim=0)


def mean_nodes(g, feat):
    x = g.ndata[feat]
    rows = []
    n0 = 0
    for n in g.node_counts():
        rows.append(x[n0:n0 + n].mean(dim=0))
        n0 += n
    return torch.stack(rows, dim=0)


def _sum_over_in_edges(g, x):
    src = torch.tensor(g.src, dtype=torch.long)
    dst = torch.tensor(g.dst, dtype=torch.long)
    out = torch.zeros((g.n,) + tuple(x.shape[1:]), dtype=x.dtype)
    if len(g.src):
        out.index_add_(0, dst, x[src])
    return out


def graph_conv(graph, feat, weight, bias, norm="both"):
    """Functional GraphConv (see the module docstring for the arithmetic)."""
    in_feats, out_feats = (weight.shape if weight is not None else (feat.shape[1], feat.shape[1]))
    if norm == "both":
        degs = graph.out_degrees().to(feat.dtype).clamp(min=1)
        feat = feat * torch.pow(degs, -0.5).reshape(-1, 1)
    if in_feats > out_feats:
        if weight is not None:
            feat = torch.matmul(feat, weight)
        rst = _sum_over_in_edges(graph, feat)
    else:
        rst = _sum_over_in_edges(graph, feat)
        if weight is not None:
            rst = torch.matmul(rst, weight)
    if norm != "none":
        degs = graph.in_degrees().to(feat.dtype).clamp(min=1)
        scale = torch.pow(degs, -0.5) if norm == "both" else 1.0 / degs
        rst = rst * scale.reshape(-1, 1)
    if bias is not None:
        rst = rst + bias
    return rst


def sage_conv_mean(graph, feat, w_self, b_self, w_neigh, b_neigh):
    """Functional SAGEConv("mean"); weights in torch.nn.Linear layout (out, in)."""
    s = _sum_over_in_edges(graph, feat)
    deg = graph.in_degrees().to(feat.dtype)
    h_neigh = torch.where(deg.reshape(-1, 1) > 0, s / deg.clamp(min=1).reshape(-1, 1), torch.zeros_like(s))
    rst = feat @ w_self.t() + h_neigh @ w_neigh.t()
    if b_self is not None:
        rst = rst + b_self
    if b_neigh is not None:
        rst = rst + b_neigh
    return rst


def gin_conv(graph, feat, eps=0.0, aggregator_type="sum"):
    """Functional GINConv body before ``apply_func`` (DGL's documented definition):
    ``(1 + eps) * x_i + aggregate_{j->i} x_j`` with aggregate = sum or mean."""
    s = _sum_over_in_edges(graph, feat)
    if aggregator_type == "mean":
        deg = graph.in_degrees().to(feat.dtype)
        s = torch.where(deg.reshape(-1, 1) > 0, s / deg.clamp(min=1).reshape(-1, 1), torch.zeros_like(s))
    return (1 + eps) * feat + s


def sym_in_powers(graph, feat, k):
    """[x, S x, ..., S^k x], S = D^-1/2 A D^-1/2 with D = in-degrees clamped to >= 1 on BOTH sides
    (DGL's SGConv / TAGConv normalisation)."""
    norm = torch.pow(graph.in_degrees().to(feat.dtype).clamp(min=1), -0.5).reshape(-1, 1)
    out = [feat]
    for _ in range(int(k)):
        out.append(_sum_over_in_edges(graph, out[-1] * norm) * norm)
    return out


def sg_conv(graph, feat, w, b, k=1):
    """Functional SGConv: ``fc(S^k x)``, ``w`` in torch.nn.Linear layout (out, in)."""
    h = sym_in_powers(graph, feat, k)[-1] @ w.t()
    return h + b if b is not None else h


def tag_conv(graph, feat, w, b, k=2):
    """Functional TAGConv: ``lin(cat([x, S x, ..., S^k x], dim=-1))``, ``w`` of shape (out, in * (k + 1))."""
    h = torch.cat(sym_in_powers(graph, feat, k), dim=-1) @ w.t()
    return h + b if b is not None else h


class GraphConv(torch.nn.Module):
    def __init__(self, in_feats, out_feats, norm="both", weight=True, bias=True, activation=None):
        super().__init__()
        self._in_feats, self._out_feats, self._norm = in_feats, out_feats, norm
        self.weight = torch.nn.Parameter(torch.Tensor(in_feats, out_feats)) if weight else None
        self.bias = torch.nn.Parameter(torch.Tensor(out_feats)) if bias else None
        self._activation = activation
        if self.weight is not None:
            torch.nn.init.xavier_uniform_(self.weight)
        if self.bias is not None:
            torch.nn.init.zeros_(self.bias)

    def forward(self, graph, feat):
        rst = graph_conv(graph, feat, self.weight, self.bias, self._norm)
        if self._activation is not None:
            rst = self._activation(rst)
        return rst


class SAGEConv(torch.nn.Module):
    def __init__(self, in_feats, out_feats, aggregator_type, feat_drop=0.0, bias=True, norm=None, activation=None):
        super().__init__()
        assert aggregator_type == "mean", "the oracle restates the 'mean' aggregator only"
        self.fc_self = torch.nn.Linear(in_feats, out_feats, bias=bias)
        self.fc_neigh = torch.nn.Linear(in_feats, out_feats, bias=bias)
        gain = torch.nn.init.calculate_gain("relu")
        torch.nn.init.xavier_uniform_(self.fc_self.weight, gain=gain)
        torch.nn.init.xavier_uniform_(self.fc_neigh.weight, gain=gain)
        self.activation, self.norm = activation, norm

    def forward(self, graph, feat):
        rst = sage_conv_mean(graph, feat, self.fc_self.weight, self.fc_self.bias,
                             self.fc_neigh.weight, self.fc_neigh.bias)
        if self.activation is not None:
            rst = self.activation(rst)
        if self.norm is not None:
            rst = self.norm(rst)
        return rst


class GINConv(torch.nn.Module):
    def __init__(self, apply_func=None, aggregator_type="sum", init_eps=0, learn_eps=False):
        super().__init__()
        self.apply_func, self._aggregator_type = apply_func, aggregator_type
        if learn_eps:
            self.eps = torch.nn.Parameter(torch.FloatTensor([init_eps]))
        else:
            self.register_buffer("eps", torch.FloatTensor([init_eps]))

    def forward(self, graph, feat):
        rst = gin_conv(graph, feat, self.eps, self._aggregator_type)
        return self.apply_func(rst) if self.apply_func is not None else rst


class SGConv(torch.nn.Module):
    def __init__(self, in_feats, out_feats, k=1, cached=False, bias=True, norm=None):
        super().__init__()
        self.fc = torch.nn.Linear(in_feats, out_feats, bias=bias)
        self._k = k
        torch.nn.init.xavier_uniform_(self.fc.weight)
        if self.fc.bias is not None:
            torch.nn.init.zeros_(self.fc.bias)

    def forward(self, graph, feat):
        return sg_conv(graph, feat, self.fc.weight, self.fc.bias, self._k)


class TAGConv(torch.nn.Module):
    def __init__(self, in_feats, out_feats, k=2, bias=True, activation=None):
        super().__init__()
        self.lin = torch.nn.Linear(in_feats * (k + 1), out_feats, bias=bias)
        self._k, self._activation = k, activation
        torch.nn.init.xavier_uniform_(self.lin.weight, gain=torch.nn.init.calculate_gain("relu"))

    def forward(self, graph, feat):
        rst = tag_conv(graph, feat, self.lin.weight, self.lin.bias, self._k)
        return self._activation(rst) if self._activation is not None else rst


def as_module(name="dgl"):
    """A module object exposing this file under DGL's names (dgl, dgl.nn.pytorch.conv)."""
    import importlib.machinery

    def _mod(n):
        m = types.ModuleType(n)
        m.__spec__ = importlib.machinery.ModuleSpec(n, None)
        m.__path__ = []
        return m

    dgl = _mod(name)
    for k in ("DGLGraph", "batch", "unbatch", "sum_nodes", "mean_nodes"):
        setattr(dgl, k, globals()[k])
    nn = _mod(name + ".nn")
    pt = _mod(name + ".nn.pytorch")
    conv = _mod(name + ".nn.pytorch.conv")
    conv.GraphConv, conv.SAGEConv = GraphConv, SAGEConv
    conv.GINConv, conv.SGConv, conv.TAGConv = GINConv, SGConv, TAGConv
    pt.conv = conv
    pt.GraphConv, pt.SAGEConv = GraphConv, SAGEConv
    nn.pytorch = pt
    dgl.nn = nn
    return dgl, nn, pt, conv


def install(name="dgl"):
    dgl, nn, pt, conv = as_module(name)
    sys.modules[name] = dgl
    sys.modules[name + ".nn"] = nn
    sys.modules[name + ".nn.pytorch"] = pt
    sys.modules[name + ".nn.pytorch.conv"] = conv
    return dgl
