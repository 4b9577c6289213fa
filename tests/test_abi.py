"""The C-ABI library loads and exports every symbol the header declares (no compute calls)."""
import ctypes
import os
import re

import pytest
import torch

from conftest import ROOT
from pinot_b200 import _lib


def _header_symbols():
    text = open(os.path.join(ROOT, "include", "pinot_b200.h")).read()
    return sorted(set(re.findall(r"PINOT_B200_API[^;(]*?\b(pinot_b200_\w+)\s*\(", text)))


def test_header_declares_the_expected_surface():
    syms = _header_symbols()
    for name in ["mp_csr_aggregate_fwd", "mp_csr_aggregate_bwd", "segment_readout_fwd", "segment_readout_bwd",
                 "node_update_fwd", "node_update_bwd", "sgemm", "gram_fwd", "gram_bwd", "potrf_ext",
                 "potrf_workspace_bytes", "trsm_rlt", "gp_nll", "gp_nll_grad", "gp_predict_mean_var", "last_error"]:
        assert "pinot_b200_" + name in syms


def test_library_exports_every_declared_symbol():
    assert os.path.exists(_lib.LIB_PATH), "run __graft_entry__.build() first"
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for name in _header_symbols():
        assert hasattr(lib, name), "missing export " + name
    # and the Python binding table covers exactly the header
    assert sorted(_lib.SIGNATURES) == _header_symbols()


def test_loader_and_pure_host_entry_points():
    lib = _lib.load()
    assert lib.pinot_b200_abi_version() == 4
    assert _lib.potrf_workspace_bytes(0) == 0
    # two column blocks: L11 + L11^-T per block, plus the (256-byte rounded) flag array
    assert _lib.potrf_workspace_bytes(65) == 2 * 2 * 64 * 64 * 4 + 256
    # workspace tail: below n = 6144 the flags of the single-launch sweep (one progress counter per 128-row tile of
    # the largest stacked matrix, 2n + r rows, plus 64 ints), from there on the hi / lo panel copies of the
    # rank-512 TMA-fed updates
    assert _lib.potrf_workspace_bytes(4200, 1) - _lib.potrf_workspace_bytes(4200) == 4 * ((8401 + 127) // 128 + 64)
    assert _lib.potrf_workspace_bytes(8192, 1) - _lib.potrf_workspace_bytes(8192) == 256 + 2 * 8193 * 512 * 4
    assert _lib.gp_nll_grad_workspace_bytes(100) == 0
    assert _lib.gp_nll_grad_workspace_bytes(4200) == 256 + 2 * 4200 * 4200 * 4
    assert _lib.last_error() == "" or isinstance(_lib.last_error(), str)


def test_cpu_tensors_are_rejected_loudly():
    """No CPU fallback: a CPU tensor must raise, never silently compute."""
    import pinot_b200 as pb
    from pinot_b200 import synthetic
    g, y = synthetic.make_batch(3, seed=1)
    net = pb.Net(pb.representation.Sequential(pb.representation.dgl_legacy.gn(), [8, "tanh"]),
                 output_regressor_class=pb.regressors.ExactGaussianProcessRegressor)
    with pytest.raises(_lib.PinotB200Error):
        net.loss(g, y)
    with pytest.raises(_lib.PinotB200Error):
        pb.regressors.kernels.RBF(4)(torch.randn(5, 4))
