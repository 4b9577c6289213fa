"""Graph convolution layers with DGL's parameter names and arithmetic, on pinot_b200 kernels.

DGL itself is not part of the reference tree (unpinned dependency); the layer
math follows DGL's documented definitions:

GraphConv(norm="both"):  h_i = in_deg_i^-1/2 * sum_{j->i} out_deg_j^-1/2 x_j, degrees clamped to >= 1,
    weight (in, out) applied BEFORE aggregation when in > out, AFTER otherwise, then + bias.
SAGEConv("mean"):        fc_self(x_i) + fc_neigh(mean_{j->i} x_j), the mean of no neighbours being 0.
GINConv("sum"):          apply_func((1 + eps) x_i + sum_{j->i} x_j), eps a buffer (0) unless learn_eps.
SGConv(k):               fc(S^k x),  S = D^-1/2 A D^-1/2 with D the IN-degrees clamped to >= 1 on both sides.
TAGConv(k):              lin([x, S x, ..., S^k x]) (concatenation along features), same S.

The last three are the other layer types pinot's ``SequentialMix`` / ``gn`` name that reduce to the same
CSR aggregate (pinot/representation/sequential.py:108-128, dgl_legacy.py:60-64).
"""
import torch

from .... import ops

__all__ = ["GraphConv", "SAGEConv", "GINConv", "SGConv", "TAGConv"]


def _csr(graph):
    if not hasattr(graph, "csr"):
        raise TypeError("expected a pinot_b200 BatchedGraph, got %r" % type(graph))
    return graph.csr()


class GraphConv(torch.nn.Module):
    def __init__(self, in_feats, out_feats, norm="both", weight=True, bias=True, activation=None,
                 allow_zero_in_degree=True):
        super().__init__()
        if norm not in ("none", "both", "right"):
            raise ValueError('Invalid norm value. Must be either "none", "both" or "right".')
        self._in_feats, self._out_feats, self._norm = in_feats, out_feats, norm
        self._allow_zero_in_degree = allow_zero_in_degree
        if weight:
            self.weight = torch.nn.Parameter(torch.empty(in_feats, out_feats))
        else:
            self.register_parameter("weight", None)
        if bias:
            self.bias = torch.nn.Parameter(torch.empty(out_feats))
        else:
            self.register_parameter("bias", None)
        self._activation = activation
        self.reset_parameters()

    def reset_parameters(self):
        if self.weight is not None:
            torch.nn.init.xavier_uniform_(self.weight)
        if self.bias is not None:
            torch.nn.init.zeros_(self.bias)

    def forward(self, graph, feat, weight=None, _act=None):
        csr = _csr(graph)
        if not self._allow_zero_in_degree and csr.has_zero_in_degree:
            from ... import DGLError
            raise DGLError("There are 0-in-degree nodes in the graph; set allow_zero_in_degree=True to "
                           "let their output be the bias only.")
        w = self.weight if weight is None else weight
        norm = {"both": "both", "none": "none", "right": "mean"}[self._norm]
        fuse = _act if self._activation is None else None
        if w is not None and self._in_feats > self._out_feats:
            # project first (cheaper), aggregate, then bias (+ activation) as an elementwise tail
            h = ops.node_update(feat, w, None, w_is_out_in=False)
            rst = ops.aggregate(h, csr, norm)
            if self.bias is not None:
                rst = rst + self.bias
            if fuse == "tanh":
                rst = torch.tanh(rst)
        else:
            rst = ops.aggregate(feat, csr, norm)
            if w is not None:
                rst = ops.node_update(rst, w, self.bias, w_is_out_in=False, act=fuse)
            else:
                if self.bias is not None:
                    rst = rst + self.bias
                if fuse == "tanh":
                    rst = torch.tanh(rst)
        if self._activation is not None:
            rst = self._activation(rst)
        return rst

    def extra_repr(self):
        return "in={_in_feats}, out={_out_feats}, normalization={_norm}".format(**self.__dict__)


class SAGEConv(torch.nn.Module):
    def __init__(self, in_feats, out_feats, aggregator_type="mean", feat_drop=0.0, bias=True, norm=None,
                 activation=None):
        super().__init__()
        if aggregator_type != "mean":
            raise NotImplementedError(
                "pinot_b200 SAGEConv implements the 'mean' aggregator (pinot's default, "
                "pinot/representation/dgl_legacy.py:18-22); got %r" % (aggregator_type,))
        self._in_feats, self._out_feats = in_feats, out_feats
        self._aggre_type = aggregator_type
        self.norm = norm
        self.feat_drop = torch.nn.Dropout(feat_drop)
        self.activation = activation
        self.fc_self = torch.nn.Linear(in_feats, out_feats, bias=bias)
        self.fc_neigh = torch.nn.Linear(in_feats, out_feats, bias=bias)
        self.reset_parameters()

    def reset_parameters(self):
        gain = torch.nn.init.calculate_gain("relu")
        torch.nn.init.xavier_uniform_(self.fc_self.weight, gain=gain)
        torch.nn.init.xavier_uniform_(self.fc_neigh.weight, gain=gain)

    def forward(self, graph, feat, _act=None):
        csr = _csr(graph)
        feat = self.feat_drop(feat)
        h_neigh = ops.aggregate(feat, csr, "mean")
        fuse = _act if (self.activation is None and self.norm is None) else None
        rst = ops.node_update(feat, self.fc_self.weight, self.fc_self.bias, w_is_out_in=True)
        rst = ops.node_update(h_neigh, self.fc_neigh.weight, self.fc_neigh.bias, y_prev=rst, w_is_out_in=True,
                              act=fuse)
        if self.activation is not None:
            rst = self.activation(rst)
        if self.norm is not None:
            rst = self.norm(rst)
        return rst


class GINConv(torch.nn.Module):
    """``apply_func((1 + eps) * x + sum_{j->i} x_j)``; ``apply_func`` is any module (pinot hands in a Linear,
    pinot/representation/dgl_legacy.py:60-64), a plain ``torch.nn.Linear`` runs on the node-update kernel."""

    def __init__(self, apply_func=None, aggregator_type="sum", init_eps=0, learn_eps=False):
        super().__init__()
        if aggregator_type not in ("sum", "mean"):
            raise NotImplementedError("pinot_b200 GINConv implements the 'sum' and 'mean' aggregators; got %r"
                                      % (aggregator_type,))
        self.apply_func = apply_func
        self._aggregator_type = aggregator_type
        if learn_eps:
            self.eps = torch.nn.Parameter(torch.FloatTensor([init_eps]))
        else:
            self.register_buffer("eps", torch.FloatTensor([init_eps]))

    def forward(self, graph, feat, _act=None):
        csr = _csr(graph)
        neigh = ops.aggregate(feat, csr, "none" if self._aggregator_type == "sum" else "mean")
        rst = (1 + self.eps) * feat + neigh
        f = self.apply_func
        if isinstance(f, torch.nn.Linear):
            return ops.node_update(rst, f.weight, f.bias, w_is_out_in=True, act=_act)
        if f is not None:
            rst = f(rst)
        return torch.tanh(rst) if _act == "tanh" else rst


def _sym_in_power(csr, feat, k):
    """[feat, S feat, ..., S^k feat] with S = D_in^-1/2 A D_in^-1/2 (degrees clamped to >= 1)."""
    out = [feat]
    for _ in range(int(k)):
        out.append(ops.aggregate(out[-1], csr, "sym_in"))
    return out


class SGConv(torch.nn.Module):
    def __init__(self, in_feats, out_feats, k=1, cached=False, bias=True, norm=None, allow_zero_in_degree=True):
        super().__init__()
        self.fc = torch.nn.Linear(in_feats, out_feats, bias=bias)
        self._k, self._cached, self._cached_h, self.norm = int(k), cached, None, norm
        self.reset_parameters()

    def reset_parameters(self):
        torch.nn.init.xavier_uniform_(self.fc.weight)
        if self.fc.bias is not None:
            torch.nn.init.zeros_(self.fc.bias)

    def forward(self, graph, feat, _act=None):
        if self._cached_h is not None:
            feat = self._cached_h
        else:
            feat = _sym_in_power(_csr(graph), feat, self._k)[-1]
            if self.norm is not None:
                feat = self.norm(feat)
            if self._cached:
                self._cached_h = feat
        return ops.node_update(feat, self.fc.weight, self.fc.bias, w_is_out_in=True, act=_act)


class TAGConv(torch.nn.Module):
    def __init__(self, in_feats, out_feats, k=2, bias=True, activation=None):
        super().__init__()
        self._in_feats, self._out_feats, self._k, self._activation = in_feats, out_feats, int(k), activation
        self.lin = torch.nn.Linear(in_feats * (self._k + 1), out_feats, bias=bias)
        self.reset_parameters()

    def reset_parameters(self):
        torch.nn.init.xavier_uniform_(self.lin.weight, gain=torch.nn.init.calculate_gain("relu"))

    def forward(self, graph, feat, _act=None):
        stack = _sym_in_power(_csr(graph), feat, self._k)
        fuse = _act if self._activation is None else None
        d = self._in_feats
        rst = None
        for i, f in enumerate(stack):            # lin(cat(stack)) as a sum of per-hop products
            last = i == len(stack) - 1
            rst = ops.node_update(f, self.lin.weight[:, i * d:(i + 1) * d], self.lin.bias if i == 0 else None,
                                  y_prev=rst, w_is_out_in=True, act=fuse if last else None)
        if self._activation is not None:
            rst = self._activation(rst)
        return rst
