"""Greedy acquisition functions of pinot_b200.active against the reference's own (pinot/active/acquisition.py),
both fed by a stand-in net whose predictive distribution is fixed (CPU)."""
import pytest
import torch

from pinot_b200 import active
from oracle import ref_loader


class _Net:
    has_exact_gp = False

    def __init__(self, mean, std):
        self.mean, self.std = mean, std

    def condition(self, gs):
        cov = torch.diag(self.std ** 2)
        return torch.distributions.MultivariateNormal(self.mean, covariance_matrix=cov)


def _net(n=40, seed=0):
    g = torch.Generator().manual_seed(seed)
    return _Net(torch.randn(n, generator=g), torch.rand(n, generator=g) + 0.1)


@pytest.mark.skipif(not ref_loader.available(), reason="reference tree not present")
@pytest.mark.parametrize("name", ["temporal", "probability_of_improvement", "uncertainty", "expected_improvement_analytical",
                                  "upper_confidence_bound"])
@pytest.mark.parametrize("q", [1, 7, 100])
def test_deterministic_acquisition_functions_match_the_reference(name, q):
    ref = ref_loader.load()
    import pinot.active.acquisition as ref_acq
    net = _net()
    y_best = torch.tensor(0.3)       # (the reference's PI needs a tensor under recent torch versions)
    want = getattr(ref_acq, name)(net, None, y_best=y_best, q=q)
    got = getattr(active, name)(net, None, y_best=y_best, q=q)
    assert torch.equal(got.cpu(), want.cpu())


@pytest.mark.skipif(not ref_loader.available(), reason="reference tree not present")
def test_sampling_acquisition_functions_match_the_reference_under_the_same_seed():
    ref_loader.load()
    import pinot.active.acquisition as ref_acq
    net = _net(60, seed=2)
    for name, kw in [("expected_improvement_monte_carlo", {"n_samples": 200}), ("random", {}), ("thompson_sampling", {})]:
        torch.manual_seed(5)
        want = getattr(ref_acq, name)(net, None, y_best=0.1, q=6, **kw)
        torch.manual_seed(5)
        got = getattr(active, name)(net, None, y_best=0.1, q=6, **kw)
        assert sorted(got.tolist()) == sorted(want.tolist()), name


def test_contract_without_the_reference():
    net = _net(10)
    for fn in (active.temporal, active.probability_of_improvement, active.uncertainty,
               active.expected_improvement_analytical, active.upper_confidence_bound, active.random,
               active.expected_improvement_monte_carlo):
        idx = fn(net, None, y_best=0.0, q=3)
        assert idx.dtype == torch.long and idx.shape == (3,) and len(set(idx.tolist())) == 3
        assert fn(net, None, q=50).shape == (10,)
    assert active.temporal(net, None, q=4).tolist() == [0, 1, 2, 3]
    assert active.uncertainty(net, None, q=1).item() == int(torch.argmax(net.std))
