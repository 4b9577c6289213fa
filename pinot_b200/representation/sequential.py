"""``pinot.representation.Sequential`` on pinot_b200 kernels.

Same constructor, attribute naming (``f_in``, ``d{idx}`` / ``a{idx}`` / ``o{idx}``, ``exes``) and
config grammar as pinot/representation/sequential.py:8-103, so reference ``state_dict``s load
unchanged.  The forward pass keeps the reference's order of operations but hands
``Linear + tanh`` (f_in) and ``conv + tanh`` pairs to one fused node-update launch each.
"""
import torch

from .. import graph as _graph
from .. import ops


def _default_pool(g):
    return _graph.sum_nodes(g, "h")


class Sequential(torch.nn.Module):
    """Sequentially staggered graph neural network (pinot/representation/sequential.py:8-103)."""

    def __init__(self, layer, config, feature_units=117, input_units=128, model_kwargs={}):
        super(Sequential, self).__init__()
        dim = input_units
        self.exes = []
        self.f_in = torch.nn.Sequential(torch.nn.Linear(feature_units, input_units), torch.nn.Tanh())
        for idx, exe in enumerate(config):
            try:
                exe = float(exe)
                if exe >= 1:
                    exe = int(exe)
            except BaseException:
                pass
            if isinstance(exe, int):      # int -> graph convolution layer
                setattr(self, "d" + str(idx), layer(dim, exe, **model_kwargs))
                dim = exe
                self.exes.append("d" + str(idx))
            elif isinstance(exe, str):    # str -> activation, a free function of torch
                setattr(self, "a" + str(idx), getattr(torch, exe))
                self.exes.append("a" + str(idx))
            elif isinstance(exe, float):  # float < 1 -> dropout
                setattr(self, "o" + str(idx), torch.nn.Dropout(exe))
                self.exes.append("o" + str(idx))

    def _node_features(self, g, x):
        if x is None:
            x = g.ndata["h"]
        lin = self.f_in[0]
        x = ops.node_update(x, lin.weight, lin.bias, w_is_out_in=True, act="tanh")
        i, n = 0, len(self.exes)
        while i < n:
            exe = self.exes[i]
            mod = getattr(self, exe)
            if exe.startswith("d"):
                nxt = getattr(self, self.exes[i + 1]) if i + 1 < n else None
                if nxt is torch.tanh and _accepts_act(mod):
                    x = mod(g, x, _act="tanh")
                    i += 2
                    continue
                x = mod(g, x)
            else:
                x = mod(x)
            i += 1
        return x

    def forward(self, g, x=None, pool=_default_pool):
        """Graph -> (batch_size, H) latent (``pool=None``: node-level (N_nodes, H))."""
        x = self._node_features(g, x)
        if pool is not None:
            with g.local_scope():
                g.ndata["h"] = x
                x = pool(g)
        return x

    def pool(self, g, h, pool=_default_pool):
        with g.local_scope():
            g.ndata["h"] = h
            return pool(g)


def _accepts_act(mod):
    import inspect
    try:
        return "_act" in inspect.signature(mod.forward).parameters
    except (TypeError, ValueError):
        return False


class SequentialMix(torch.nn.Module):
    """Mixed-layer variant (pinot/representation/sequential.py:257-366): the config is a flat list of
    strings, a layer-type keyword followed by its parameters in the order of the reference's
    ``layer_param_type`` table (:108-128), e.g. ``["GraphConv", "32", "activation", "tanh", "TAGConv", "64",
    "2", "dropout", "0.1"]``.  Layer types on the CSR kernels: GraphConv, SAGEConv (mean), GINConv (Linear
    apply function, sum), SGConv, TAGConv; the others the reference's table names (GATConv, EdgeConv,
    RelGraphConv, GatedGraphConv, ChebConv) need per-edge attention / relation kernels that are outside
    SURVEY.md section 8 and raise NotImplementedError."""

    LAYER_PARAMS = {"GraphConv": (), "SAGEConv": (), "GINConv": (), "SGConv": ("k",), "TAGConv": ("k",)}

    def __init__(self, config, feature_units=117, input_units=128, model_kwargs={}, input_transform=True):
        super(SequentialMix, self).__init__()
        from ..dgl.nn.pytorch import conv as _conv
        last = feature_units
        self.input_transform = input_transform
        if input_transform:
            last = input_units
            self.f_in = torch.nn.Sequential(torch.nn.Linear(feature_units, input_units), torch.nn.Tanh())
        self.exes = []
        i, idx = 0, 0
        config = list(config)
        while i < len(config):
            name = config[i]
            if name in self.LAYER_PARAMS:
                out = int(config[i + 1])
                extra = self.LAYER_PARAMS[name]
                kw = {k: int(config[i + 2 + j]) for j, k in enumerate(extra)}
                if name == "SAGEConv":
                    kw["aggregator_type"] = "mean"
                if name == "GINConv":
                    layer = _conv.GINConv(apply_func=torch.nn.Linear(last, out), aggregator_type="sum")
                else:
                    layer = getattr(_conv, name)(last, out, **kw)
                setattr(self, "conv" + str(idx), layer)
                self.exes.append("conv" + str(idx))
                last = out
                i += 2 + len(extra)
            elif name == "activation":
                setattr(self, "act" + str(idx), getattr(torch, config[i + 1]))
                self.exes.append("act" + str(idx))
                i += 2
            elif name == "dropout":
                setattr(self, "drop" + str(idx), torch.nn.Dropout(float(config[i + 1])))
                self.exes.append("drop" + str(idx))
                i += 2
            else:
                raise NotImplementedError(
                    "SequentialMix layer type %r is outside the B200 hot path (available: %s, activation, dropout)"
                    % (name, ", ".join(self.LAYER_PARAMS)))
            idx += 1
        self.dummy_output = torch.nn.Linear(1, last)

    def forward(self, g, h=None, pool=_default_pool):
        if h is None:
            h = g.ndata["h"]
        if hasattr(self, "f_in"):
            lin = self.f_in[0]
            h = ops.node_update(h, lin.weight, lin.bias, w_is_out_in=True, act="tanh")
        with g.local_scope():
            for exe in self.exes:
                mod = getattr(self, exe)
                h = mod(g, h) if exe.startswith("conv") else mod(h)
            if pool is not None:
                g.ndata["h"] = h
                h = pool(g)
            return h

    def pool(self, g, h, pool=_default_pool):
        with g.local_scope():
            g.ndata["h"] = h
            return pool(g)
