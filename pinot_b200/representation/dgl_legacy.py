"""``pinot.representation.dgl_legacy`` on pinot_b200 kernels (pinot/representation/dgl_legacy.py:15-72).

``GN`` wraps a graph-convolution layer and exposes ``in_features`` / ``out_features`` so that
``pinot.Net`` can discover the representation width (pinot/net.py:141-145); ``gn(model_name)`` is
the layer factory handed to ``Sequential``.
"""
import torch

from ..dgl.nn.pytorch import conv as _conv

DEFAULT_MODEL_KWARGS = {
    "SAGEConv": {"aggregator_type": "mean"},
    "GATConv": {"num_heads": 3},
    "TAGConv": {"k": 2},
}


class GN(torch.nn.Module):
    """Graph net layer; state_dict keys are ``gn.weight`` / ``gn.bias`` (GraphConv) or
    ``gn.fc_self.*`` / ``gn.fc_neigh.*`` (SAGEConv), as in the reference."""

    def __init__(self, in_features, out_features, model_name="GraphConv", kwargs={}):
        super(GN, self).__init__()
        if kwargs == {}:
            if model_name in DEFAULT_MODEL_KWARGS:
                kwargs = DEFAULT_MODEL_KWARGS[model_name]
        if not hasattr(_conv, model_name):
            raise NotImplementedError(
                "pinot_b200 implements GraphConv and SAGEConv (SURVEY.md section 8a); %r is not available"
                % (model_name,))
        self.gn = getattr(_conv, model_name)(in_features, out_features, **kwargs)
        self.in_features = in_features
        self.out_features = out_features

    def forward(self, g, x, _act=None):
        if _act is not None:
            return self.gn(g, x, _act=_act)
        return self.gn(g, x)


def gn(model_name="GraphConv", kwargs={}):
    """Layer factory: ``gn()(in_features, out_features) -> GN``."""
    return lambda in_features, out_features: GN(
        in_features=in_features, out_features=out_features, model_name=model_name, kwargs=kwargs,
    )
