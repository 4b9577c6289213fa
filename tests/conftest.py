import os
import sys
import warnings

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")

warnings.filterwarnings("ignore", category=UserWarning)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def load_golden(name):
    return torch.load(os.path.join(GOLDEN, name + ".pt"), map_location="cpu", weights_only=False)


NET_CASES = ["net_gcn32_exact_rbf", "net_gcn128_exact_rbf", "net_gcn64to32_exact_rbf", "net_sage128_exact_matern",
             "net_sage64_vgp"]
GP_EXACT_CASES = ["gp_exact_rbf_n300", "gp_exact_rbf_n131_d7", "gp_exact_matern_n150"]
GP_VGP_CASES = ["gp_vgp_rbf_n200_m64"]


def rel_err(a, b, nan_equal=False):
    a = a.detach().double().cpu().reshape(-1)
    b = b.detach().double().cpu().reshape(-1)
    if nan_equal:
        # the reference itself produces NaN in places (Matern52 gradients): require the same pattern
        if not torch.equal(torch.isnan(a), torch.isnan(b)):
            return float("inf")
        keep = ~torch.isnan(b)
        a, b = a[keep], b[keep]
        if a.numel() == 0:
            return 0.0
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def max_rel_err(a, b, floor=1e-6):
    """max |a-b| / max(|b|, floor * max|b|): element-wise relative error with a sane floor."""
    a = a.detach().double().cpu().reshape(-1)
    b = b.detach().double().cpu().reshape(-1)
    scale = b.abs().clamp_min(floor * float(b.abs().max().clamp_min(1e-30)))
    return float(((a - b).abs() / scale).max())
