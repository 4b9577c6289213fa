"""``pinot.Net`` for the B200 kernels: a representation module feeding a regressor head.

Contract taken from the reference (pinot/net.py:15-248) and pinned by tests/test_api.py:

* ``Net(representation, output_regressor_class=NeuralNetworkRegressor, output_regressor=None,
  output_likelihood_class=None, **head_kwargs)``; attributes ``representation``,
  ``output_regressor``, ``representation_out_features``, ``has_exact_gp``, and -- after a ``loss`` call
  with an exact-GP head -- ``g_last`` / ``y_last``;
* the head's width is read from the LAST sub-module of the representation that exposes
  ``out_features`` (pinot/net.py:141-145), so any ``nn.Module`` can be a representation
  (pinot/active/tests use ``Linear + Tanh``);
* ``loss(g, y)`` embeds and delegates to the head's own ``loss`` when it has one, else to the negative
  log-probability of ``condition`` (pinot/net.py:59-98); with an exact GP it also remembers the batch,
  because ``condition(g)`` conditions the posterior on the embeddings of that batch under the CURRENT
  weights (pinot/net.py:191-193, :240-242).

Beyond the reference: the training-set embeddings and their factorisation are cached between
``condition`` / ``condition_marginal`` calls.  The cache key is the identity of
``g_last`` / ``y_last`` plus the version counter of every parameter, so an optimizer step, a
``load_state_dict`` or a new ``loss`` call invalidates it and the reference's semantics are kept.
"""
import torch

from .regressors import ExactGaussianProcessRegressor, NeuralNetworkRegressor


def _last_out_features(module):
    width = None
    for sub in module.modules():
        if hasattr(sub, "out_features"):
            width = sub.out_features
    if width is None:
        raise ValueError("the representation has no sub-module with an `out_features` attribute")
    return width


class _PosteriorCache:
    """Holder that never travels with ``copy.deepcopy`` / pickling of the net (it can be gigabytes)."""

    def __init__(self):
        self.key, self.h_last, self.posterior = None, None, None

    def clear(self):
        self.key, self.h_last, self.posterior = None, None, None

    def __deepcopy__(self, memo):
        return _PosteriorCache()

    def __reduce__(self):
        return (_PosteriorCache, ())


class BaseNet(torch.nn.Module):
    """Representation + head; subclasses provide ``condition``."""

    def __init__(self, representation, output_regressor_class=None, *args, **kwargs):
        super().__init__()
        self.representation = representation

    def condition(self, g, *args, sampler=None, **kwargs):
        raise NotImplementedError

    def _condition(self, h, *args, **kwargs):
        return self.output_regressor.condition(h, *args, **kwargs)

    def _loss(self, h, y, *args, **kwargs):
        head_loss = getattr(self.output_regressor, "loss", None)
        if head_loss is not None:
            return head_loss(h, y, *args, **kwargs)
        return -self._condition(h, *args, **kwargs).log_prob(y).sum()

    def loss(self, g, y, *args, **kwargs):
        return self._loss(self.representation(g), y, *args, **kwargs)


class Net(BaseNet):
    def __init__(self, representation, output_regressor_class=NeuralNetworkRegressor, output_regressor=None,
                 output_likelihood_class=None, **kwargs):
        super().__init__(representation)
        self.representation_out_features = _last_out_features(representation)
        if output_regressor is None and output_regressor_class is not None:
            output_regressor = output_regressor_class(in_features=self.representation_out_features, **kwargs)
        self.has_exact_gp = isinstance(output_regressor, ExactGaussianProcessRegressor)
        if output_likelihood_class is not None:
            output_regressor = output_likelihood_class(base_regressor=output_regressor)
        self.output_regressor = output_regressor
        self._cache = _PosteriorCache()

    # -- training ------------------------------------------------------------------------------
    def loss(self, g, y, *args, **kwargs):
        if self.has_exact_gp:
            self.g_last, self.y_last = g, y
        return super().loss(g, y, *args, **kwargs)

    # -- prediction ----------------------------------------------------------------------------
    def _cache_key(self):
        stochastic = self.training and any(
            isinstance(m, torch.nn.Dropout) and m.p > 0 for m in self.representation.modules())
        if stochastic:
            return None
        g, y = self.g_last, self.y_last
        feats = g.ndata["h"] if hasattr(g, "ndata") and "h" in g.ndata else (g if torch.is_tensor(g) else None)
        return (id(g), id(y), None if feats is None else (feats.data_ptr(), feats._version), y._version,
                tuple((p.data_ptr(), p._version) for p in self.parameters()))

    def _training_embeddings(self):
        """``representation(g_last)`` under the current weights, computed once per (batch, weights)."""
        key = self._cache_key()
        if key is None or not hasattr(self.g_last, "csr"):
            return self.representation(self.g_last), None
        if self._cache.key != key:
            self._cache.clear()
            with torch.no_grad():
                self._cache.h_last = self.representation(self.g_last)
            self._cache.key = key
        return self._cache.h_last, self._cache

    def _posterior(self):
        """Cached training factorisation (ops.ExactGPPosterior) for the exact-GP head."""
        from . import ops
        reg = self.output_regressor
        h_last, cache = self._training_embeddings()
        if cache is not None and cache.posterior is not None:
            return cache.posterior
        y_last = self.y_last
        if hasattr(reg, "_gathered"):      # sharded head: condition on every rank's training rows
            h_last, y_last = reg._gathered(h_last, y_last)
        with torch.no_grad():
            post = ops.ExactGPPosterior(h_last, y_last, reg.kernel.scale, reg.kernel.variance, reg.log_sigma,
                                        kind=reg._kind(), check=bool(reg.check_positive_definite))
        if cache is not None:
            cache.posterior = post
        return post

    def condition(self, g, *args, sampler=None, **kwargs):
        h = self.representation(g)
        sampling = sampler is not None and hasattr(sampler, "sample_params")
        if self.has_exact_gp:
            h_last, cache = self._training_embeddings()      # both embeddings precede the parameter draw
        if sampling:
            sampler.sample_params()
            cache = None                                     # the head's parameters just changed
        if self.has_exact_gp:
            head = self.output_regressor
            if (cache is not None and not args and "x_tr" not in kwargs and "y_tr" not in kwargs
                    and hasattr(head, "condition_from_posterior") and not hasattr(head, "_gathered")):
                # same batch, same weights as an earlier call: reuse its factorisation
                return head.condition_from_posterior(self._posterior(), h)
            kwargs.setdefault("x_tr", h_last)
            kwargs.setdefault("y_tr", self.y_last)
        return self._condition(h, *args, **kwargs)

    def condition_marginal(self, g, chunk_molecules=65536, row_offset=0):
        """Per-molecule predictive ``Normal`` for an exact-GP head without the n_te x n_te
        covariance (not in the reference API; equals ``pinot.metrics._independent(net.condition(g))``).
        The training graphs are embedded and factored once per (batch, weights); the graphs in ``g`` are
        embedded and scored ``chunk_molecules`` at a time, so million-molecule pools fit.  ``row_offset`` is the index of
        ``g``'s first molecule in the whole test set when ``g`` is one shard of it (it places the ``1e-5`` that
        ``_perturb`` puts on the main diagonal of the rectangular K_te,tr, gaussian_process_regressor.py:95)."""
        if not self.has_exact_gp:
            raise NotImplementedError("condition_marginal is available for exact-GP heads only")
        with torch.no_grad():
            post = self._posterior()
            nb = g.batch_size
            means, variances = [], []
            for lo in range(0, nb, chunk_molecules):
                hi = min(nb, lo + chunk_molecules)
                sub = g if (lo == 0 and hi == nb) else g.select(range(lo, hi))
                m, v = post.mean_var(self.representation(sub), row_offset=int(row_offset) + lo)
                means.append(m)
                variances.append(v)
        mean = torch.cat(means) if len(means) > 1 else means[0]
        var = torch.cat(variances) if len(variances) > 1 else variances[0]
        return torch.distributions.normal.Normal(loc=mean, scale=var.clamp_min(0).sqrt(), validate_args=False)
