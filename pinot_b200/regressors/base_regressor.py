"""Base class of the output heads (pinot/regressors/base_regressor.py:11-27)."""
import abc

import torch


class BaseRegressor(torch.nn.Module):
    def __init__(self):
        super(BaseRegressor, self).__init__()

    @abc.abstractmethod
    def condition(self, h, *args, **kwargs):
        raise NotImplementedError
