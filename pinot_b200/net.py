"""``pinot.Net``: representation -> regressor glue (pinot/net.py:15-248), unchanged semantics.

* ``representation_out_features`` is read from the last sub-module exposing ``out_features``
  (:141-145); any ``nn.Module`` works as representation (pinot/active/tests use Linear + Tanh);
* with an exact-GP head ``loss`` remembers the last batch (``g_last``, ``y_last``, :191-193) and
  ``condition`` re-embeds it to condition the posterior on it (:240-242).
"""
import abc

import torch

from .regressors import NeuralNetworkRegressor, ExactGaussianProcessRegressor


class BaseNet(torch.nn.Module, abc.ABC):
    def __init__(self, representation, output_regressor_class, *args, **kwargs):
        super(BaseNet, self).__init__()
        self.representation = representation

    @abc.abstractmethod
    def condition(self, g, sampler=None, *args, **kwargs):
        raise NotImplementedError

    def loss(self, g, y, *args, **kwargs):
        h = self.representation(g)
        return self._loss(h, y, *args, **kwargs)

    def _loss(self, h, y, *args, **kwargs):
        # use the regressor's own loss when it has one (:93-94)
        if hasattr(self.output_regressor, "loss"):
            return self.output_regressor.loss(h, y, *args, **kwargs)
        distribution = self._condition(h, *args, **kwargs)
        nll = -distribution.log_prob(y).sum()
        return nll


class Net(BaseNet):
    def __init__(self, representation, output_regressor_class=NeuralNetworkRegressor, output_regressor=None,
                 output_likelihood_class=None, **kwargs):
        super(Net, self).__init__(
            representation=representation,
            output_regressor_class=output_regressor_class,
            output_likelihood_class=output_likelihood_class,
        )
        self.representation_out_features = [
            layer for layer in list(self.representation.modules()) if hasattr(layer, "out_features")
        ][-1].out_features

        if output_regressor is None and output_regressor_class is not None:
            output_regressor = output_regressor_class(in_features=self.representation_out_features, **kwargs)

        self.has_exact_gp = False
        if isinstance(output_regressor, ExactGaussianProcessRegressor):
            self.has_exact_gp = True

        self.representation = representation

        if output_likelihood_class is not None:
            self.output_regressor = output_likelihood_class(base_regressor=output_regressor)
        else:
            self.output_regressor = output_regressor

    def loss(self, g, y, *args, **kwargs):
        h = self.representation(g)
        if self.has_exact_gp is True:
            self.g_last = g
            self.y_last = y
        return self._loss(h, y, *args, **kwargs)

    def _condition(self, h, *args, **kwargs):
        return self.output_regressor.condition(h, *args, **kwargs)

    def condition(self, g, *args, **kwargs):
        h = self.representation(g)
        sampler = None
        if "sampler" in kwargs:
            sampler = kwargs.pop("sampler")
        if self.has_exact_gp is True:
            h_last = self.representation(self.g_last)
            kwargs = {**{"x_tr": h_last, "y_tr": self.y_last}, **kwargs}
        if sampler is not None and hasattr(sampler, "sample_params"):
            sampler.sample_params()
        return self._condition(h, *args, **kwargs)

    def condition_marginal(self, g, chunk_molecules=65536):
        """Per-molecule predictive ``Normal`` for an exact-GP head without the n_te x n_te
        covariance (not in the reference API; equals ``pinot.metrics._independent(net.condition(g))``).
        The training graphs are embedded and factored once; the graphs in ``g`` are embedded and
        scored ``chunk_molecules`` at a time, so million-molecule pools fit."""
        if not self.has_exact_gp:
            raise NotImplementedError("condition_marginal is available for exact-GP heads only")
        from . import ops
        reg = self.output_regressor
        with torch.no_grad():
            h_last, y_last = self.representation(self.g_last), self.y_last
            if hasattr(reg, "_gathered"):      # sharded head: condition on every rank's training rows
                h_last, y_last = reg._gathered(h_last, y_last)
            post = ops.ExactGPPosterior(h_last, y_last, reg.kernel.scale, reg.kernel.variance, reg.log_sigma,
                                        kind=reg._kind(), check=bool(reg.check_positive_definite))
            nb = g.batch_size
            means, variances = [], []
            for lo in range(0, nb, chunk_molecules):
                hi = min(nb, lo + chunk_molecules)
                sub = g if (lo == 0 and hi == nb) else g.select(range(lo, hi))
                m, v = post.mean_var(self.representation(sub))
                means.append(m)
                variances.append(v)
        mean = torch.cat(means) if len(means) > 1 else means[0]
        var = torch.cat(variances) if len(variances) > 1 else variances[0]
        return torch.distributions.normal.Normal(loc=mean, scale=var.clamp_min(0).sqrt(), validate_args=False)
