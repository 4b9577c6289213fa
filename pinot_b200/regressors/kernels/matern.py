"""Matern-5/2 kernel on the B200 Gram kernel (pinot/regressors/kernels/matern.py:10-63):
``exp(variance) * (1 + sqrt(5 d^2) + 5/3 d^2) * exp(-sqrt(5 d^2))`` with a SCALAR log length-scale."""
import torch

from ... import ops
from .base_kernel import BaseKernel


class Matern52(BaseKernel):
    kind = "matern52"

    def __init__(self, scale=0.0, variance=0.0):
        # Same signature as the reference.  NB the regressors build their kernel as
        # ``kernel(in_features)`` (gaussian_process_regressor.py:79), so ``kernel=Matern52`` makes
        # ``scale = in_features`` there, exactly as in the reference; pass
        # ``kernel=lambda in_features: Matern52()`` for a unit length-scale.
        super(Matern52, self).__init__()
        self.scale = torch.nn.Parameter(torch.tensor(float(scale)))
        self.variance = torch.nn.Parameter(torch.tensor(float(variance)))

    def forward(self, x, x_=None):
        assert x.dim() == 2
        if x_ is not None:
            assert x_.dim() == 2
            if x_ is x:
                x_ = None
        return ops.gram(self.kind, x, x_, self.scale, self.variance)
