"""The scratch-free sweep entry point (``pinot_b200_potrf_ext_noalloc``) against the regular one.

It exists for hosts that can pass buffers to this library but cannot get new device memory when
they call it (a plug-in inside a runtime that owns the GPU's memory pool, a loop that may not
trigger cudaMalloc's device-wide synchronisation).  The two entry points compute the same
factorisation on different schedules (this one: the plain two-launches-per-panel sweep with its
deferred diagonal blocks parked in the library's built-in array), so their outputs must agree; two calls
in a row (a large problem, then a small one) check that contents left in the built-in array by one
call cannot leak into the next.  Collected last on purpose (file name), and self-contained so that
it also passes alone in a fresh process.
"""
import pytest
import torch

from pinot_b200 import _lib
from pinot_b200._lib import call, ptr

pytestmark = pytest.mark.gpu
DEV = "cuda"
MAX_N = 8192  # PINOT_B200_POTRF_NOALLOC_MAX_N in include/pinot_b200.h


def _problem(n, r, seed):
    g = torch.Generator().manual_seed(seed)
    a = torch.randn(n, n, generator=g, dtype=torch.float64)
    k = a @ a.t() / n + 0.5 * torch.eye(n, dtype=torch.float64)
    ld = (n + 3) // 4 * 4
    e = torch.zeros(n + r + n, ld, dtype=torch.float32)
    e[:n, :n] = k.float()
    e[n:n + r, :n] = torch.randn(r, n, generator=g)
    e[n + r:, :n].diagonal().fill_(1.0)          # identity rows -> L^-T
    return e.to(DEV), ld


def _run(entry, e, ld, n, r):
    e = e.clone()
    diag = torch.empty(n, device=DEV)
    info = torch.full((1,), -7, dtype=torch.int32, device=DEV)
    if entry == "potrf_ext":
        ws_bytes = _lib.potrf_workspace_bytes(n)
        ws = torch.empty(max(ws_bytes // 4, 1), device=DEV)
        call(entry, ptr(e), ld, n, r, 1, ptr(diag), ptr(ws), ws_bytes, ptr(info))
    else:
        call(entry, ptr(e), ld, n, r, 1, ptr(diag), ptr(info))
    torch.cuda.synchronize()
    return e, diag, int(info.item())


def _rel(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm())


def test_noalloc_sweep_matches_the_workspace_entry_point_twice_in_a_row():
    for n, r, seed in [(1000, 3, 1), (130, 2, 2)]:   # second call is smaller: stale scratch must not matter
        e, ld = _problem(n, r, seed)
        want, want_diag, want_info = _run("potrf_ext", e, ld, n, r)
        got, got_diag, got_info = _run("potrf_ext_noalloc", e, ld, n, r)
        assert want_info == 0 and got_info == 0
        assert _rel(torch.tril(got[:n, :n]), torch.tril(want[:n, :n])) < 1e-5      # L
        assert _rel(got[n:], want[n:]) < 1e-5                                      # extension rows, L^-T
        assert _rel(got_diag, want_diag) < 1e-5
        # and the factor is right in its own terms: L L^T reproduces the input
        l = torch.tril(got[:n, :n]).double()
        assert _rel(l @ l.t(), torch.tril(e[:n, :n]).double() + torch.tril(e[:n, :n], -1).double().t()) < 1e-5


def test_noalloc_sweep_rejects_sizes_beyond_its_built_in_scratch():
    n = MAX_N + 1
    e = torch.empty(1, device=DEV)               # never touched: the size check comes first
    info = torch.zeros(1, dtype=torch.int32, device=DEV)
    with pytest.raises(_lib.PinotB200Error, match="exceeds the built-in scratch limit of %d" % MAX_N):
        call("potrf_ext_noalloc", ptr(e), n, n, 0, 0, None, ptr(info))
    # a non-positive-definite input is reported through info exactly like the regular entry point
    bad = torch.eye(70, device=DEV)
    bad[40, 40] = -1.0
    stacked = torch.cat([bad, torch.eye(70, device=DEV)]).contiguous()      # [A ; I], ld = 70
    _, _, code = _run("potrf_ext_noalloc", stacked, 70, 70, 0)
    assert code == 41
