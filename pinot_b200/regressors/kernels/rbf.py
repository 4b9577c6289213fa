"""pinot's "RBF" kernel on the B200 Gram kernel (pinot/regressors/kernels/rbf.py:10-55).

NB it is ``exp(variance) * exp(-0.5 * ||(x - x') * exp(scale)||_2)`` -- the UN-squared Euclidean
distance -- with ARD length-scales ``scale`` (in_features,) by default.
"""
import torch

from ... import ops
from .base_kernel import BaseKernel


class RBF(BaseKernel):
    kind = "rbf"

    def __init__(self, in_features, scale=0.0, variance=0.0, ard=True):
        super(RBF, self).__init__()
        if ard is True:
            self.scale = torch.nn.Parameter(scale * torch.ones(in_features))
        else:
            self.scale = torch.nn.Parameter(torch.tensor(scale))
        self.variance = torch.nn.Parameter(torch.tensor(variance))

    def forward(self, x, x_=None):
        """(n, d), (n', d) -> (n, n') covariance; differentiable in x, x_, scale, variance."""
        assert x.dim() == 2
        if x_ is not None:
            assert x_.dim() == 2
            if x_ is x:
                x_ = None
        return ops.gram(self.kind, x, x_, self.scale, self.variance)
