"""Fast data plumbing and acquisition for pinot's Bayesian-optimisation loop.

``pinot.active.experiment.BayesOptExperiment`` takes ``slice_fn`` / ``collate_fn`` / ``acquisition``
as arguments; passing the functions below keeps its loop unchanged while
  * re-batching a selection of molecules is index arithmetic plus a row-gather kernel on the device
    instead of ``dgl.unbatch`` + ``dgl.batch`` over Python lists (pinot/active/experiment.py:83-129),
  * Thompson sampling reads per-molecule marginals (``Net.condition_marginal``) instead of building an
    n_pool x n_pool covariance and factoring it (pinot/active/acquisition.py:72-118).
"""
import torch
import torch.distributed as dist

from . import graph as _graph


def slice_fn_tensor(x, idxs):
    """pinot/active/experiment.py:29-46."""
    return x[torch.as_tensor(list(idxs), dtype=torch.long, device=x.device)]


def collate_fn_tensor(x):
    return torch.stack(x)


def collate_fn_graph(x):
    return _graph.batch(x)


def slice_fn_graph(x, idxs):
    """pinot/active/experiment.py:83-104 with the same error behaviour."""
    if x.batch_size > 1:
        return x.select(idxs)
    raise RuntimeError("Can only slice if there is more than one.")


def slice_fn_tuple(x, idxs):
    """pinot/active/experiment.py:107-129."""
    gs, ys = x
    idxs = list(idxs)
    return slice_fn_graph(gs, idxs), slice_fn_tensor(ys, idxs)


def _independent(net, gs, row_offset=0):
    """``pinot.metrics._independent(net.condition(gs))`` -- through the marginal path when the net
    offers one.  ``row_offset``: pool index of the first molecule of ``gs`` when it is one shard of the pool."""
    if hasattr(net, "condition_marginal") and getattr(net, "has_exact_gp", False):
        return net.condition_marginal(gs, row_offset=row_offset)
    d = net.condition(gs)
    return torch.distributions.normal.Normal(loc=d.mean.flatten(), scale=d.variance.pow(0.5).flatten())


def thompson_sampling(net, gs, y_best=0.0, q=1, unique=True):
    """Same contract as pinot/active/acquisition.py:72-118: ``q`` posterior draws, the arg-max of each,
    duplicates replaced by further draws when ``unique``.  Returns a LongTensor of pool indices."""
    distribution = _independent(net, gs)
    thetas = distribution.sample((q,))
    pending_pts = torch.argmax(thetas, axis=1)
    if unique:
        pending_pts = set(torch.unique(pending_pts).tolist())
        while len(pending_pts) < q:
            theta = distribution.sample()
            pending_pts.add(torch.argmax(theta).item())
        pending_pts = list(pending_pts)
    return torch.as_tensor(pending_pts, dtype=torch.long).cpu()


def thompson_sampling_sharded(net, gs_local, global_offset, q=1, generator=None):
    """Thompson sampling over a pool that is block-partitioned across ranks:
    every rank scores its own molecules and draws ``q`` posterior samples over them; the global
    arg-max of each draw is found with one all-gather of (value, index) pairs.  ``global_offset`` is
    the pool index of this rank's first molecule.  Returns ``q`` global indices (duplicates possible),
    identical on every rank."""
    distribution = _independent(net, gs_local, row_offset=int(global_offset))
    n_local = distribution.loc.numel()
    eps = torch.randn((q, n_local), device=distribution.loc.device, generator=generator)
    thetas = distribution.loc + distribution.scale * eps
    val, idx = thetas.max(dim=1)
    idx = idx + int(global_offset)
    if dist.is_initialized() and dist.get_world_size() > 1:
        w = dist.get_world_size()
        vals = [torch.empty_like(val) for _ in range(w)]
        idxs = [torch.empty_like(idx) for _ in range(w)]
        dist.all_gather(vals, val)
        dist.all_gather(idxs, idx)
        vals, idxs = torch.stack(vals), torch.stack(idxs)          # (world, q)
        best = vals.argmax(dim=0)
        idx = idxs.gather(0, best[None, :])[0]
    return idx.cpu()


# ------------------------------------------------------------------------------------------------
# greedy acquisition functions over the predictive marginals (pinot/active/acquisition.py:10-69,121-400)
# ------------------------------------------------------------------------------------------------
def _top(utility, q):
    """Indices of the ``q`` largest utilities (all of them if there are fewer), as in ``_greedy`` (:21-27)."""
    return torch.topk(utility, min(q, len(utility))).indices


def temporal(net, gs, y_best=0.0, q=1):
    """The first ``q`` candidates in data-set order (:29-36, :121-161; baseline for temporal data sets)."""
    d = _independent(net, gs)
    n = d.mean.flatten().shape[0]
    return _top(torch.arange(n - 1, -1, -1, dtype=torch.float32, device=d.mean.device), q)


def probability_of_improvement(net, gs, y_best=0.0, q=1):
    """P(y > y_best) under the predictive marginal (:38-39, :164-201)."""
    d = _independent(net, gs)
    # (a float y_best is accepted: recent torch versions insist on a tensor argument to cdf)
    return _top(1.0 - d.cdf(torch.as_tensor(y_best, dtype=d.mean.dtype, device=d.mean.device)), q)


def uncertainty(net, gs, y_best=0.0, q=1):
    """Predictive variance (:41-42, :204-241)."""
    return _top(_independent(net, gs).variance, q)


def expected_improvement_analytical(net, gs, y_best=0.0, q=1):
    """Closed-form expected improvement of a normal marginal (:44-53, :244-292)."""
    d = _independent(net, gs)
    mu, sigma = d.mean, d.stddev
    z = (mu - y_best) / sigma
    unit = torch.distributions.Normal(0, 1)
    return _top((mu - y_best) * unit.cdf(z) + sigma * torch.exp(unit.log_prob(z)), q)


def expected_improvement_monte_carlo(net, gs, y_best=0.0, q=1, n_samples=1000):
    """Sample mean of max(y - y_best, 0) (:55-59, :295-334)."""
    d = _independent(net, gs)
    return _top(torch.nn.functional.relu(d.sample((n_samples,)) - y_best).mean(axis=0), q)


def upper_confidence_bound(net, gs, y_best=0.0, q=1, kappa=0.95):
    """Upper end of the central ``kappa`` interval (:61-64, :337-373; pinot/samplers/utils.py:62-84)."""
    d = _independent(net, gs)
    return _top(d.icdf(1 - (1 - torch.tensor(kappa)) / 2), q)


def random(net, gs, y_best=0.0, q=1, seed=2666):
    """Uniform random utilities (:66-67, :376-414; like the reference, ``seed`` is accepted and not used)."""
    d = _independent(net, gs)
    return _top(torch.rand(d.batch_shape), q)
