"""Layer factory of ``pinot.representation.dgl_legacy`` on pinot_b200 kernels.

Contract (pinot/representation/dgl_legacy.py:15-72, pinned by tests/test_api.py):

* ``gn(model_name="GraphConv", kwargs={})`` returns ``layer(in_features, out_features)``, the factory
  that ``Sequential`` calls once per integer entry of its config;
* the layer it builds is a ``GN``: attribute ``gn`` holds the convolution (so state_dict keys read
  ``d{idx}.gn.weight``, ``d{idx}.gn.fc_self.weight`` ...), and ``in_features`` / ``out_features`` are
  exposed for ``Net`` to read the representation width from (pinot/net.py:141-145);
* with no ``kwargs`` a layer type takes the defaults of ``DEFAULT_MODEL_KWARGS`` (SAGEConv -> mean
  aggregator, TAGConv -> k = 2);
* ``gn("GINConv")`` is the one exception: it returns the bare GINConv with a ``Linear`` apply function
  and the sum aggregator (:60-64).
"""
import torch

from ..dgl.nn.pytorch import conv as _conv

DEFAULT_MODEL_KWARGS = {
    "SAGEConv": {"aggregator_type": "mean"},
    "GATConv": {"num_heads": 3},
    "TAGConv": {"k": 2},
}


def _conv_class(model_name):
    cls = getattr(_conv, model_name, None) if model_name in _conv.__all__ else None
    if cls is None:
        raise NotImplementedError(
            "pinot_b200 implements %s on its CSR kernels (SURVEY.md sections 8a, 8f-4); %r is not available"
            % (", ".join(_conv.__all__), model_name))
    return cls


class GN(torch.nn.Module):
    """One graph-convolution layer with the width attributes pinot expects."""

    def __init__(self, in_features, out_features, model_name="GraphConv", kwargs={}):
        super().__init__()
        options = dict(kwargs) if kwargs else dict(DEFAULT_MODEL_KWARGS.get(model_name, {}))
        self.in_features, self.out_features = in_features, out_features
        self.gn = _conv_class(model_name)(in_features, out_features, **options)

    def forward(self, g, x, _act=None):
        # `_act` lets Sequential hand the following tanh to the layer's fused epilogue
        return self.gn(g, x) if _act is None else self.gn(g, x, _act=_act)


def gn(model_name="GraphConv", kwargs={}):
    """``gn(name)(in_features, out_features)`` -> layer module."""
    if model_name == "GINConv":
        def make_gin(in_features, out_features):
            return _conv.GINConv(apply_func=torch.nn.Linear(in_features, out_features), aggregator_type="sum")
        return make_gin

    def make(in_features, out_features):
        return GN(in_features, out_features, model_name=model_name, kwargs=kwargs)
    return make
