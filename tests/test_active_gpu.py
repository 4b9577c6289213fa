"""Active-learning plumbing on the GPU: device-side re-batching, chunked marginal posterior,
Thompson sampling, and a miniature Bayesian-optimisation loop in the shape of
pinot/active/experiment.py:361-402 (blind pick -> train -> acquire -> update)."""
import random

import pytest
import torch

import pinot_b200 as pb
from pinot_b200 import active, ops, synthetic
from conftest import rel_err
from oracle import gp_ref

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _net(hidden=32):
    torch.manual_seed(0)
    net = pb.Net(pb.representation.Sequential(pb.representation.dgl_legacy.gn(), [hidden, "tanh", hidden, "tanh"]),
                 output_regressor_class=pb.regressors.ExactGaussianProcessRegressor).to(DEV)
    with torch.no_grad():
        net.output_regressor.kernel.scale.fill_(-2.0)
    return net


def test_select_on_device_matches_host():
    g, _ = synthetic.make_batch(200, seed=21)
    idxs = [199, 0, 57, 57, 3, 120]
    want = g.select(idxs)
    got = g.to(DEV).select(idxs)
    assert torch.equal(got.edges()[0].cpu(), want.edges()[0]) and torch.equal(got.edges()[1].cpu(), want.edges()[1])
    assert torch.equal(got.ndata["h"].cpu(), want.ndata["h"])
    assert got.batch_num_nodes().tolist() == want.batch_num_nodes().tolist()


def test_posterior_cache_matches_oracle_and_chunking_is_exact():
    torch.manual_seed(3)
    n, nte, d = 300, 700, 32            # more test than training points: the rectangular perturb hits every chunk rule
    x, xt = torch.randn(n, d) * 0.5, torch.randn(nte, d) * 0.5
    y = torch.randn(n, 1)
    sc, var, ls = torch.full((d,), -0.5), torch.tensor(0.3), torch.tensor(-2.0)
    mean_w, cov_w = gp_ref.exact_gp_condition(xt.double(), x.double(), y.double(), sc.double(), var.double(), ls.double())
    post = ops.ExactGPPosterior(x.to(DEV), y.to(DEV), sc.to(DEV), var.to(DEV), ls.to(DEV))
    mean, v = post.mean_var(xt.to(DEV))
    assert rel_err(mean, mean_w) < 1e-4 and rel_err(v, torch.diagonal(cov_w)) < 1e-4
    mean_c, v_c = post.mean_var(xt.to(DEV), chunk=128)
    # chunks below 256 test points take the CUDA-core Gram route, the unchunked call the tensor-core one: equal to
    # the rounding of the two routes (each within 1e-4 of the reference), and the rectangular perturb lands on the
    # same global diagonal either way
    assert rel_err(mean_c, mean) < 2e-5 and rel_err(v_c, v) < 2e-5
    mean_d, v_d = post.mean_var(xt.to(DEV), chunk=350)      # both chunkings on the tensor-core route
    assert rel_err(mean_d, mean) < 1e-6 and rel_err(v_d, v) < 1e-6


def test_condition_marginal_chunked_equals_full_condition():
    net = _net()
    g, y = synthetic.make_batch(64, seed=5, device=DEV)
    pool, _ = synthetic.make_batch(150, seed=6, device=DEV)
    net.loss(g, y)
    full = net.condition(pool)
    marg = net.condition_marginal(pool, chunk_molecules=40)
    # two algebraically equal fp32 routes (K_te,tr alpha with alpha = L^-T z, versus (K_te,tr L^-T) z):
    # they agree to a few cond(K) * eps, not bit for bit
    assert rel_err(marg.mean, full.mean) < 5e-4
    assert rel_err(marg.variance, full.variance) < 1e-3


def test_thompson_sampling_contract():
    net = _net()
    g, y = synthetic.make_batch(48, seed=7, device=DEV)
    pool, _ = synthetic.make_batch(300, seed=8, device=DEV)
    net.loss(g, y)
    pts = active.thompson_sampling(net, pool, q=10)
    assert pts.dtype == torch.long and pts.numel() == 10 and len(set(pts.tolist())) == 10
    assert int(pts.min()) >= 0 and int(pts.max()) < 300
    pts2 = active.thompson_sampling_sharded(net, pool, global_offset=1000, q=5)
    assert pts2.numel() == 5 and int(pts2.min()) >= 1000 and int(pts2.max()) < 1300


def test_miniature_bayesian_optimisation_loop():
    """blind_pick -> (train -> acquire -> update_data) x rounds, as BayesOptExperiment.run does."""
    net = _net()
    data = synthetic.make_batch(120, seed=9, device=DEV)
    opt = torch.optim.Adam(net.parameters(), 1e-3)
    random.seed(2666)
    unseen = list(range(120))
    seen = [unseen.pop(random.choice(unseen))]
    q = 6
    for _ in range(3):
        seen_data = active.slice_fn_tuple(data, seen) if len(seen) > 1 else (data[0].select(seen), data[1][seen])
        unseen_data = active.slice_fn_tuple(data, unseen)
        for _epoch in range(2):
            opt.zero_grad()
            loss = net.loss(*seen_data).sum()
            loss.backward()
            opt.step()
        assert torch.isfinite(loss)
        pending = active.thompson_sampling(net, unseen_data[0], y_best=float(seen_data[1].max()), q=q)
        pending = pending.sort(descending=True).values
        seen.extend([unseen.pop(int(p)) for p in pending])
    assert len(seen) == 1 + 3 * q and len(set(seen)) == len(seen) and len(unseen) == 120 - len(seen)
