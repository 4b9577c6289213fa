"""Default head of ``Net``: a Normal likelihood whose parameters are linear in the embedding.

Outside the GP hot path (component 4); plain PyTorch, present so that
``pinot_b200.Net(representation)`` without a head argument behaves like the reference's
(pinot/regressors/neural_network_regressor.py:11-110): ``condition(h)`` returns
``Normal(mu(h), sigma)`` where sigma is ``exp`` of a second linear output ("normal-heteroschedastic",
the default), ``exp`` of one free parameter ("normal-homoschedastic") or fixed at 1
("normal-homoschedastic-fixed").  There is no ``loss``: ``Net`` falls back to ``-log_prob(y).sum()``.
"""
import torch

from .base_regressor import BaseRegressor

_NOISE_MODELS = ("normal-heteroschedastic", "normal-homoschedastic", "normal-homoschedastic-fixed")


class NeuralNetworkRegressor(BaseRegressor):
    def __init__(self, in_features, output_regression=None, noise_model="normal-heteroschedastic",
                 measurement_dimension=1):
        super().__init__()
        self.noise_model = noise_model
        self.measurement_dimension = measurement_dimension
        if output_regression is None:
            # two linear read-outs: location and (for the heteroschedastic model) log-scale
            self._output_regression = torch.nn.ModuleList(
                torch.nn.Linear(in_features, measurement_dimension) for _ in range(2))
            output_regression = self._linear_readouts
        self.output_regression = output_regression

    def _linear_readouts(self, theta):
        return [head(theta) for head in self._output_regression]

    def _scale(self, mu, log_sigma):
        shape = (1, self.measurement_dimension)
        if self.noise_model == "normal-heteroschedastic":
            return torch.exp(log_sigma)
        if self.noise_model == "normal-homoschedastic":
            if getattr(self, "LOG_SIGMA", None) is None:      # created on first use, on the data's device
                self.LOG_SIGMA = torch.zeros(shape, device=mu.device, requires_grad=True)
            return torch.exp(self.LOG_SIGMA)
        if self.noise_model == "normal-homoschedastic-fixed":
            return torch.ones(shape, device=mu.device)
        raise ValueError("unknown noise model %r (expected one of %s)" % (self.noise_model, ", ".join(_NOISE_MODELS)))

    def condition(self, h, **kwargs):
        mu, log_sigma = self.output_regression(h)
        return torch.distributions.normal.Normal(loc=mu, scale=self._scale(mu, log_sigma))
