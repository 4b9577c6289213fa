"""Maximum-likelihood Normal head (pinot/regressors/neural_network_regressor.py:11-110).

Not part of the GP hot path (component 4): kept in plain PyTorch so that
``pinot_b200.Net``'s default ``output_regressor_class`` behaves like the reference's.
"""
import torch

from .base_regressor import BaseRegressor


class NeuralNetworkRegressor(BaseRegressor):
    def __init__(self, in_features, output_regression=None, noise_model="normal-heteroschedastic",
                 measurement_dimension=1):
        super(NeuralNetworkRegressor, self).__init__()
        if output_regression is None:
            self._output_regression = torch.nn.ModuleList(
                [torch.nn.Linear(in_features, measurement_dimension) for _ in range(2)]
            )

            def output_regression(theta):
                return [f(theta) for f in self._output_regression]

        self.noise_model = noise_model
        self.measurement_dimension = measurement_dimension
        self.output_regression = output_regression

    def condition(self, h, **kwargs):
        theta = self.output_regression(h)
        if self.noise_model == "normal-heteroschedastic":
            mu, log_sigma = theta
            return torch.distributions.normal.Normal(loc=mu, scale=torch.exp(log_sigma))
        if self.noise_model == "normal-homoschedastic":
            mu, _ = theta
            if not hasattr(self, "LOG_SIGMA"):
                self.LOG_SIGMA = torch.zeros((1, self.measurement_dimension), device=mu.device)
                self.LOG_SIGMA.requires_grad = True
            return torch.distributions.normal.Normal(loc=mu, scale=torch.exp(self.LOG_SIGMA))
        if self.noise_model == "normal-homoschedastic-fixed":
            mu, _ = theta
            return torch.distributions.normal.Normal(
                loc=mu, scale=torch.ones((1, self.measurement_dimension), device=mu.device))
        raise ValueError("unknown noise model %r" % (self.noise_model,))
