"""Gaussian-process heads on pinot_b200 kernels.

Same classes, constructor arguments, parameter names and return types as
pinot/regressors/gaussian_process_regressor.py:17-508.  Reference quirks that parity depends on are
kept on purpose: the exact-GP objective uses ``trace(L)`` where the log
marginal likelihood has ``sum(log(diag(L)))`` (:160-164); ``_perturb`` adds ``1e-5`` to the main
diagonal of every kernel matrix, rectangular ones included (:29-46, :95); the exact predictive
covariance carries no observation noise (:223-226) while the variational one does (:416-418).
"""
import math

import numpy as np
import torch

from .. import ops
from .base_regressor import BaseRegressor
from .kernels.rbf import RBF


class GaussianProcessRegressor(BaseRegressor):
    """Gaussian Process Regression (base)."""

    def __init__(self, epsilon=1e-5):
        super(GaussianProcessRegressor, self).__init__()
        self.epsilon = epsilon

    def condition(self, *args, **kwargs):
        raise NotImplementedError

    def _perturb(self, k):
        """``k + epsilon * eye(*k.shape)`` (:29-46); the fused paths fold this into the Gram launch."""
        noise = self.epsilon * torch.eye(*k.shape, device=k.device)
        return k + noise


def _mvn_from_cov(mean, cov, check=True):
    """``MultivariateNormal(mean, cov)`` whose internal Cholesky factor comes from the library's own
    sweep instead of torch.linalg (the reference lets torch factor ``cov`` inside the constructor,
    gaussian_process_regressor.py:229-231)."""
    scale_tril = ops.cholesky(cov, check=check)
    return torch.distributions.multivariate_normal.MultivariateNormal(
        loc=mean, scale_tril=scale_tril, validate_args=False)


class ExactGaussianProcessRegressor(GaussianProcessRegressor):
    """Exact GP head (gaussian_process_regressor.py:52-233).

    Extra, optional keyword (not in the reference): ``logdet="trace"`` (default, the reference's
    objective) or ``"exact"`` (true negative log marginal likelihood).
    """

    def __init__(self, in_features, kernel=None, log_sigma=-3.0, logdet="trace"):
        super(ExactGaussianProcessRegressor, self).__init__()
        if kernel is None:
            kernel = RBF
        self.kernel_cls = kernel
        self.kernel = kernel(in_features)
        self.in_features = in_features
        self.log_sigma = torch.nn.Parameter(torch.tensor(log_sigma))
        assert logdet in ("trace", "exact")
        self.logdet = logdet
        #: True: raise RuntimeError on a non-positive-definite kernel matrix right away, like
        #: torch.cholesky (one device->host read in the middle of the step); "deferred": keep the
        #: factorisation status on the device and raise from verify(), which the caller runs where it
        #: synchronises anyway (after reading the loss); False: never check.
        self.check_positive_definite = True
        self._last_info = None

    def _kind(self):
        kind = getattr(self.kernel, "kind", None)
        if kind is None:
            raise NotImplementedError(
                "pinot_b200 GP heads need a kernel with a fused Gram implementation (RBF, Matern52)")
        return kind

    def _get_kernel_and_auxiliary_variables(self, x_tr, y_tr, x_te=None):
        """Unfused restatement of :85-121 (kept for API parity; ``loss`` / ``condition`` use the
        fused sweep).  Returns (k_tr_tr, k_te_te, k_te_tr, k_tr_te, l_low, alpha)."""
        k_tr_tr = ops.gram(self._kind(), x_tr, None, self.kernel.scale, self.kernel.variance, self.epsilon)
        if x_te is not None:
            k_te_te = ops.gram(self._kind(), x_te, None, self.kernel.scale, self.kernel.variance, self.epsilon)
            k_te_tr = ops.gram(self._kind(), x_te, x_tr, self.kernel.scale, self.kernel.variance, self.epsilon)
            k_tr_te = k_te_tr.t()
        else:
            k_te_te = k_te_tr = k_tr_te = k_tr_tr
        n = k_tr_tr.shape[0]
        k_plus_sigma = k_tr_tr + torch.exp(self.log_sigma) * torch.eye(n, device=k_tr_tr.device)
        # one sweep with right-hand sides [y | I]: L^-1 y and L^-1, hence alpha = L^-T (L^-1 y)
        rhs = torch.cat([y_tr.reshape(n, -1), torch.eye(n, device=k_tr_tr.device)], dim=1)
        l_low, sol = ops.chol_solve(k_plus_sigma, rhs, check=(self.check_positive_definite is True))
        ny = rhs.shape[1] - n
        alpha = ops.matmul(sol[:, ny:], sol[:, :ny], True, False)
        return k_tr_tr, k_te_te, k_te_tr, k_tr_te, l_low, alpha

    def loss(self, x_tr, y_tr, *args, **kwargs):
        """Negative log likelihood, shape (1, 1) (:123-166)."""
        self._x_tr = x_tr
        self._y_tr = y_tr
        immediate = self.check_positive_definite is True
        loss, self._last_info = ops.exact_gp_loss(
            x_tr, y_tr, self.kernel.scale, self.kernel.variance, self.log_sigma, kind=self._kind(),
            exact_logdet=(self.logdet == "exact"), check=immediate, return_info=True)
        return loss

    def verify(self):
        """Raise RuntimeError if the last ``loss`` factorisation met a non-positive pivot (for
        ``check_positive_definite="deferred"``).  Reads one int from the device."""
        if self._last_info is not None and self.check_positive_definite:
            ops.check_info(self._last_info)

    def condition(self, x_te, *args, x_tr=None, y_tr=None, **kwargs):
        """Predictive distribution ``MultivariateNormal(mean, cov)`` (:168-233).  Evaluated without
        autograd history (acquisition functions and metrics only consume values)."""
        mean, cov = ops.exact_gp_posterior(
            x_te, x_tr, y_tr, self.kernel.scale, self.kernel.variance, self.log_sigma, kind=self._kind(),
            full_cov=True, check=bool(self.check_positive_definite))
        return _mvn_from_cov(mean, cov, check=bool(self.check_positive_definite))

    def condition_from_posterior(self, posterior, x_te):
        """``condition`` from a training factorisation that is already available (``ops.ExactGPPosterior``,
        kept by ``Net`` between calls): no second sweep over the training set."""
        mean, cov = posterior.mean_cov(x_te)
        return _mvn_from_cov(mean, cov, check=bool(self.check_positive_definite))

    def condition_marginal(self, x_te, x_tr=None, y_tr=None):
        """Predictive marginals ``Normal(mean, sqrt(var))`` without materialising the n_te x n_te
        covariance: what ``pinot.metrics._independent`` (pinot/metrics.py:13-27) extracts from
        ``condition`` and all that Thompson sampling consumes (pinot/active/acquisition.py:72-118).
        Not in the reference API."""
        mean, var = ops.exact_gp_posterior(
            x_te, x_tr, y_tr, self.kernel.scale, self.kernel.variance, self.log_sigma, kind=self._kind(),
            full_cov=False, check=bool(self.check_positive_definite))
        return torch.distributions.normal.Normal(loc=mean, scale=var.clamp_min(0).sqrt(), validate_args=False)


class ShardedExactGaussianProcessRegressor(ExactGaussianProcessRegressor):
    """Exact GP head over the UNION of all ranks' training rows (not in the reference, which is
    single-device; arithmetic of gaussian_process_regressor.py:85-233 on the gathered rows).

    ``loss(x_tr_local, y_tr_local)`` takes THIS rank's rows (the embeddings of its molecules), runs the
    Gram matrix and the Cholesky sweep sharded by column blocks over the process group
    (pinot_b200.sharded_gp) and returns the same (1, 1) value on every rank.  ``backward`` gives each rank
    the gradient of its own rows; the kernel hyper-parameters and ``log_sigma`` receive per-rank partial
    sums, so the data-parallel gradient all-reduce (``GradientBucket.allreduce(average=False)``) completes
    them -- do not average.  ``condition`` / ``condition_marginal`` gather the training rows and predict
    this rank's test rows with the replicated single-GPU posterior.
    """

    def __init__(self, in_features, kernel=None, log_sigma=-3.0, logdet="trace", group=None, block=512):
        super(ShardedExactGaussianProcessRegressor, self).__init__(in_features, kernel=kernel, log_sigma=log_sigma,
                                                                   logdet=logdet)
        self.group = group
        self.block = block

    def loss(self, x_tr, y_tr, *args, **kwargs):
        from .. import sharded_gp
        self._x_tr = x_tr
        self._y_tr = y_tr
        loss = sharded_gp.sharded_exact_gp_loss(
            x_tr, y_tr, self.kernel.scale, self.kernel.variance, self.log_sigma, kind=self._kind(),
            exact_logdet=(self.logdet == "exact"), group=self.group, block=self.block,
            check=(self.check_positive_definite is True))
        self._last_info = sharded_gp._ShardedExactGPLoss.last_info
        return loss

    def _gathered(self, x_tr, y_tr):
        from .. import sharded_gp
        import torch.distributed as dist
        world = dist.get_world_size(self.group) if (dist.is_available() and dist.is_initialized()) else 1
        x_all, _ = sharded_gp._gather_rows(x_tr.detach().contiguous(), self.group, world)
        y_all, _ = sharded_gp._gather_rows(y_tr.detach().reshape(x_tr.shape[0], -1).contiguous(), self.group, world)
        return x_all, y_all

    def condition(self, x_te, *args, x_tr=None, y_tr=None, **kwargs):
        x_all, y_all = self._gathered(x_tr, y_tr)
        return super(ShardedExactGaussianProcessRegressor, self).condition(x_te, x_tr=x_all, y_tr=y_all)

    def condition_marginal(self, x_te, x_tr=None, y_tr=None):
        x_all, y_all = self._gathered(x_tr, y_tr)
        return super(ShardedExactGaussianProcessRegressor, self).condition_marginal(x_te, x_tr=x_all, y_tr=y_all)


class VariationalGaussianProcessRegressor(GaussianProcessRegressor):
    """Sparse variational GP head (gaussian_process_regressor.py:236-508)."""

    def __init__(self, in_features, kernel=None, log_sigma=-3.0, n_inducing_points=100, mu_initializer_std=0.1,
                 sigma_initializer_value=-2.0, kl_loss_scaling=1e-2, grid_boundary=1.0):
        super(VariationalGaussianProcessRegressor, self).__init__()
        if kernel is None:
            kernel = RBF
        self.kernel_cls = kernel
        self.kernel = kernel(in_features)
        self.in_features = in_features
        self.log_sigma = torch.nn.Parameter(torch.tensor(log_sigma))
        # inducing inputs ~ U(-b, b), (M, H) (:280-291)
        self.x_tr = torch.nn.Parameter(
            torch.distributions.uniform.Uniform(
                -1.0 * grid_boundary * torch.ones(n_inducing_points, in_features),
                1.0 * grid_boundary * torch.ones(n_inducing_points, in_features),
            ).sample())
        # variational mean ~ N(0, std^2), (M, 1) (:293-300)
        self.y_tr_mu = torch.nn.Parameter(
            torch.distributions.normal.Normal(
                loc=torch.zeros(n_inducing_points, 1),
                scale=mu_initializer_std * torch.ones(n_inducing_points, 1),
            ).sample())
        self.y_tr_sigma_diag = torch.nn.Parameter(sigma_initializer_value * torch.ones(n_inducing_points))
        self.y_tr_sigma_tril = torch.nn.Parameter(
            torch.zeros(int(n_inducing_points * (n_inducing_points - 1) * 0.5)))
        self.n_inducing_points = n_inducing_points
        self.kl_loss_scaling = kl_loss_scaling
        #: True / "deferred" / False, as for the exact head: "deferred" keeps the status of every
        #: factorisation of the last ``loss`` / ``forward`` on the device until ``verify()`` reads it
        self.check_positive_definite = True
        self._infos = []

    def verify(self):
        """Raise RuntimeError if a factorisation of the last ``loss`` / ``forward`` call met a non-positive
        pivot (for ``check_positive_definite="deferred"``).  One device read per factorisation."""
        if self.check_positive_definite:
            for info in self._infos:
                ops.check_info(info)

    def _kind(self):
        kind = getattr(self.kernel, "kind", None)
        if kind is None:
            raise NotImplementedError("pinot_b200 GP heads need an RBF or Matern52 kernel")
        return kind

    def _y_tr_sigma(self):
        """Lower-triangular variational scale: exp(diag) on the diagonal, ``y_tr_sigma_tril`` filled
        row-major below it (:316-332)."""
        m = self.y_tr_sigma_diag.shape[0]
        y_tr_diag = torch.diag_embed(torch.exp(self.y_tr_sigma_diag))
        # The reference fills the strictly-lower part with masked_scatter (row-major order of the mask).  The same
        # positions as explicit indices: masked_scatter and its backward (masked_select) read a count back from the
        # device, which stalls the step and cannot be captured in a CUDA graph.
        key = (m, y_tr_diag.device)
        if getattr(self, "_tril_index_key", None) != key:
            rows, cols = torch.tril_indices(m, m, offset=-1, device=y_tr_diag.device)      # row-major, like the mask
            self._tril_index, self._tril_index_key = (rows, cols), key
        return y_tr_diag.index_put(self._tril_index, self.y_tr_sigma_tril)

    def _gram(self, a, b=None):
        return ops.gram(self._kind(), a, b, self.kernel.scale, self.kernel.variance, self.epsilon)

    def _k_tr_tr(self):
        return self._gram(self.x_tr)

    def _k_tr_te(self, x_te):
        return self._gram(self.x_tr, x_te)

    def _k_te_tr(self, x_te):
        return self._k_tr_te(x_te).t()

    def _k_te_te(self, x_te):
        return self._gram(x_te)

    def _forward_parts(self, x_te):
        """One sweep over ``[K_uu ; mu^T ; S^T ; K_fu]`` gives L, L^-1 mu, L^-1 S and L^-1 K_uf
        (:373-390)."""
        m = self.n_inducing_points
        k_tr_tr = self._k_tr_tr()
        k_tr_te = self._k_tr_te(x_te)
        rhs = torch.cat([self.y_tr_mu, self._y_tr_sigma(), k_tr_te], dim=1)
        self._infos = []
        l, sol = ops.chol_solve(k_tr_tr, rhs, check=(self.check_positive_definite is True), info_sink=self._infos)
        return l, sol[:, :1], sol[:, 1:1 + m], sol[:, 1 + m:]

    def forward(self, x_te):
        """(n, H) -> predictive mean (n, 1) and covariance (n, n) (:346-420)."""
        predictive_mean, predictive_cov, _ = self._forward(x_te)
        return predictive_mean, predictive_cov

    def _forward(self, x_te):
        l, linv_mu, linv_sigma, a = self._forward_parts(x_te)
        n = x_te.shape[0]
        predictive_mean = ops.matmul(a, linv_mu, True, False)          # A^T L^-1 mu
        b = ops.matmul(linv_sigma, a, True, False)                      # (L^-1 S)^T A
        predictive_cov = (
            self._k_te_te(x_te)
            - ops.matmul(a, a, True, False)
            + ops.matmul(b, b, True, False)
        )
        predictive_cov = predictive_cov + torch.exp(self.log_sigma) * torch.eye(n, device=x_te.device)
        return predictive_mean, predictive_cov, (l, linv_mu, linv_sigma)

    def condition(self, x_te, *args, **kwargs):
        predictive_mean, predictive_cov = self.forward(x_te)
        scale_tril = ops.cholesky(predictive_cov, check=(self.check_positive_definite is True))
        return torch.distributions.multivariate_normal.MultivariateNormal(
            loc=predictive_mean.flatten(), scale_tril=scale_tril, validate_args=False)

    def loss(self, x_te, y_te, *args, **kwargs):
        """``-log N(y | m, cov).mean() + kl_loss_scaling * (log q(u) - E_q log p(u))`` (:446-486);
        the result is float64 like the reference's."""
        predictive_mean, predictive_cov, (l, linv_mu, linv_sigma) = self._forward(x_te)
        n = x_te.shape[0]
        m = self.n_inducing_points
        resid = (y_te.flatten() - predictive_mean.flatten()).reshape(n, 1)
        l_cov, r = ops.chol_solve(predictive_cov, resid, check=(self.check_positive_definite is True),
                                  info_sink=self._infos)
        half_log_det = torch.log(torch.diagonal(l_cov)).sum()
        log_prob = -0.5 * (r * r).sum() - half_log_det - 0.5 * n * math.log(2.0 * math.pi)
        nll = -log_prob  # `.mean()` of a scalar (:473)

        # E_q[log p(u)] with p = N(0, K_uu), q = N(mu, S S^T) (:488-502).  The reference repeats the
        # triangular solves in float64; L^-1 S and L^-1 mu are already available from the sweep and
        # only the reductions are carried out in float64 here.
        const_term = -0.5 * m * np.log(2 * np.pi * np.exp(1))
        log_det_prior = -torch.log(torch.diagonal(l)).double().sum()
        trace_term = -0.5 * (linv_sigma.double() ** 2).sum()
        quad_term = -0.5 * (linv_mu.double() ** 2).sum()
        log_p_u = const_term + log_det_prior + trace_term + quad_term
        # -entropy of q (:504-508): log det S = sum(y_tr_sigma_diag)
        log_q_u = -(self.y_tr_sigma_diag.sum() + 0.5 * m * np.log(2 * np.pi * np.exp(1)))
        return nll + self.kl_loss_scaling * (log_q_u - log_p_u)

    @staticmethod
    def exp_log_full_gaussian(mean1, tril1, mean2, tril2):
        """E_{N(mean1, tril1 tril1^T)}[log N(. | mean2, tril2 tril2^T)] (:488-502)."""
        m = mean1.shape[0]
        const_term = -0.5 * m * np.log(2 * np.pi * np.exp(1))
        log_det_prior = -torch.sum(torch.log(torch.diag(tril2)))
        rhs = torch.cat([tril1, mean1 - mean2], dim=1)
        _, sol = ops.chol_solve(ops.matmul(tril2, tril2, False, True), rhs)
        trace_term = -0.5 * torch.sum(sol[:, :m].double() ** 2)
        quad_term = -0.5 * torch.sum(sol[:, m:].double() ** 2)
        return const_term + log_det_prior + trace_term + quad_term

    @staticmethod
    def entropy_full_gaussian(mean1, tril1):
        const_term = 0.5 * mean1.shape[0] * np.log(2 * np.pi * np.exp(1))
        log_det_prior = torch.sum(torch.log(torch.diag(tril1)))
        return log_det_prior + const_term
