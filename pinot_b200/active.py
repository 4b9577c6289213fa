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


def _independent(net, gs):
    """``pinot.metrics._independent(net.condition(gs))`` -- through the marginal path when the net
    offers one."""
    if hasattr(net, "condition_marginal") and getattr(net, "has_exact_gp", False):
        return net.condition_marginal(gs)
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
    distribution = _independent(net, gs_local)
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
