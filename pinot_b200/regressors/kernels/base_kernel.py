"""Base class of the GP covariance functions (pinot/regressors/kernels/base_kernel.py:10-32)."""
import abc

import torch


class BaseKernel(torch.nn.Module, abc.ABC):
    """A Gaussian-process kernel that hosts its hyper-parameters."""

    #: name understood by the C-ABI Gram kernels ("rbf" | "matern52")
    kind = None

    def __init__(self, epsilon=1e-6):
        super(BaseKernel, self).__init__()
        self.epsilon = epsilon

    @abc.abstractmethod
    def forward(self, x, *args, **kwargs):
        raise NotImplementedError
