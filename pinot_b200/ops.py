"""torch.autograd wrappers around the C-ABI kernels (``include/pinot_b200.h``).

PyTorch is used for device memory, streams and autograd bookkeeping only; every O(N*H), O(N^2) or
O(N^3) operation below is one of the library's own kernels.  All tensors must live on a CUDA
device: there is no CPU path.
"""
import math

import torch

from . import _lib
from ._lib import call, ptr

EPSILON = 1e-5  # GaussianProcessRegressor.epsilon, pinot/regressors/gaussian_process_regressor.py:20
KIND = {"rbf": 0, "matern52": 1}
#: the exact-GP loss obtains L^-T from N identity extension rows riding through the sweep (bench.py counts
#: their N^3/3 flops as algorithmic work of the potrf_ext call)
LOSS_SWEEP_IDENTITY_ROWS = True


def _f32c(t):
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


def _pad4(n):
    return (int(n) + 3) // 4 * 4


# ------------------------------------------------------------------------------------------------
# kernel (1): message passing + readout
# ------------------------------------------------------------------------------------------------
class _Aggregate(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, csr, norm):
        x = _f32c(x)
        src_scale, dst_scale = csr.scales(norm)
        out = torch.empty_like(x)
        call("mp_csr_aggregate_fwd", ptr(x), ptr(out), ptr(csr.indptr), ptr(csr.indices), ptr(src_scale),
             ptr(dst_scale), x.shape[0], x.shape[1])
        ctx.csr, ctx.norm = csr, norm
        return out

    @staticmethod
    def backward(ctx, dout):
        dout = _f32c(dout)
        csr = ctx.csr
        src_scale, dst_scale = csr.scales(ctx.norm)
        dx = torch.empty_like(dout)
        call("mp_csr_aggregate_bwd", ptr(dout), ptr(dx), ptr(csr.indptr_t), ptr(csr.indices_t), ptr(src_scale),
             ptr(dst_scale), dout.shape[0], dout.shape[1])
        return dx, None, None


def aggregate(x, csr, norm="both"):
    """``D_in^-1/2 A D_out^-1/2 x`` (norm="both"), ``D_in^-1 A x`` ("mean") or ``A x`` ("none")."""
    return _Aggregate.apply(x, csr, norm)


class _SegmentReadout(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, csr, mean):
        x = _f32c(x)
        out = torch.empty((csr.batch_size, x.shape[1]), device=x.device, dtype=torch.float32)
        call("segment_readout_fwd", ptr(x), ptr(out), ptr(csr.seg_offsets), csr.batch_size, x.shape[1], int(mean))
        ctx.csr, ctx.mean, ctx.n = csr, int(mean), x.shape[0]
        return out

    @staticmethod
    def backward(ctx, dout):
        dout = _f32c(dout)
        csr = ctx.csr
        dx = torch.empty((ctx.n, dout.shape[1]), device=dout.device, dtype=torch.float32)
        call("segment_readout_bwd", ptr(dout), ptr(dx), ptr(csr.seg_offsets), ptr(csr.node_seg), ctx.n,
             dout.shape[1], ctx.mean)
        return dx, None, None


def segment_readout(x, csr, mean=False):
    return _SegmentReadout.apply(x, csr, mean)


# ------------------------------------------------------------------------------------------------
# kernel (2): node update
# ------------------------------------------------------------------------------------------------
class _NodeUpdate(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, b, y_prev, w_is_out_in, act):
        x, w = _f32c(x), _f32c(w)
        n, fin = x.shape
        fout = w.shape[0] if w_is_out_in else w.shape[1]
        assert (w.shape[1] if w_is_out_in else w.shape[0]) == fin, "node_update: weight/in_features mismatch"
        bb = _f32c(b) if b is not None else None
        if y_prev is not None:
            y = _f32c(y_prev).clone()
        else:
            y = torch.empty((n, fout), device=x.device, dtype=torch.float32)
        call("node_update_fwd", ptr(x), ptr(w), ptr(bb), ptr(y), n, fin, fout, int(w_is_out_in), int(act),
             int(y_prev is not None))
        ctx.save_for_backward(x, w, y if act else None)
        ctx.cfg = (int(w_is_out_in), int(act), b is not None, y_prev is not None)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w, y = ctx.saved_tensors
        w_is_out_in, act, has_b, has_prev = ctx.cfg
        dy = _f32c(dy)
        n, fin = x.shape
        fout = dy.shape[1]
        need_x, need_w, need_b, need_prev = ctx.needs_input_grad[:4]
        dz = torch.empty_like(dy) if act else None
        dx = torch.empty_like(x) if need_x else None
        dw = torch.empty_like(w) if need_w else None
        db = torch.empty((fout,), device=dy.device, dtype=torch.float32) if (has_b and need_b) else None
        call("node_update_bwd", ptr(dy), ptr(y), ptr(x), ptr(w), ptr(dz), ptr(dx), ptr(dw), ptr(db), n, fin, fout,
             w_is_out_in, act)
        dprev = (dz if act else dy) if (has_prev and need_prev) else None
        return dx, dw, db, dprev, None, None


def node_update(x, w, b=None, y_prev=None, w_is_out_in=True, act=None):
    """``act(x W + b [+ y_prev])``; ``act`` is None or "tanh"."""
    assert act in (None, "tanh"), "fused activation must be None or 'tanh'"
    base = getattr(x, "_pb_pitch_base", None)
    if (base is not None and not x.requires_grad and x.dtype == torch.float32 and x.dim() == 2 and x.data_ptr() == base.data_ptr()
            and base.shape[0] == x.shape[0] and x.stride() == (base.shape[1], 1)):
        # x is the (n, F) view of a zero-padded (n, pad4(F)) buffer (expand_atom_features(pad_pitch=True)): multiply the
        # whole buffer by the weight with zero rows appended -- the same sum, on 16-byte aligned rows
        pad = base.shape[1] - x.shape[1]
        w = torch.nn.functional.pad(w, (0, pad) if w_is_out_in else (0, 0, 0, pad))
        x = base
    return _NodeUpdate.apply(x, w, b, y_prev, bool(w_is_out_in), 1 if act == "tanh" else 0)


class _MatMul(torch.autograd.Function):
    @staticmethod
    def forward(ctx, a, b, ta, tb):
        a, b = _f32c(a), _f32c(b)
        m = a.shape[1] if ta else a.shape[0]
        k = a.shape[0] if ta else a.shape[1]
        n = b.shape[0] if tb else b.shape[1]
        assert (b.shape[1] if tb else b.shape[0]) == k, "matmul: inner dimensions differ"
        c = torch.empty((m, n), device=a.device, dtype=torch.float32)
        call("sgemm", int(ta), int(tb), m, n, k, 1.0, ptr(a), a.shape[1], ptr(b), b.shape[1], 0.0, ptr(c), n)
        ctx.save_for_backward(a, b)
        ctx.t = (ta, tb)
        return c

    @staticmethod
    def backward(ctx, dc):
        a, b = ctx.saved_tensors
        ta, tb = ctx.t
        da = db = None
        if ctx.needs_input_grad[0]:
            # C = op(A) op(B): d op(A) = dC op(B)^T
            da = matmul(dc, b, False, not tb) if not ta else matmul(b, dc, tb, True)
        if ctx.needs_input_grad[1]:
            db = matmul(a, dc, not ta, False) if not tb else matmul(dc, a, True, ta)
        return da, db, None, None


def matmul(a, b, ta=False, tb=False):
    """``op(a) @ op(b)`` on the library's fp32 GEMM (differentiable)."""
    return _MatMul.apply(a, b, bool(ta), bool(tb))


# ------------------------------------------------------------------------------------------------
# kernel (3): Gram matrix
# ------------------------------------------------------------------------------------------------
def gram_raw(kind, x, y, scale, variance, out=None, ldo=None, diag_add=0.0, diag_add_log=None, lower_only=False):
    """Non-differentiable Gram launch; ``out`` may be a view into a larger row-major buffer.
    ``lower_only``: skip the tiles strictly above the diagonal (the sweep reads the lower triangle only)."""
    n, d = x.shape
    n2 = y.shape[0]
    if out is None:
        out = torch.empty((n, n2), device=x.device, dtype=torch.float32)
        ldo = n2
    same = x.data_ptr() == y.data_ptr() and n == n2
    ws_bytes = _lib.gram_workspace_bytes(n, n2, d, same, False)        # 0: small problem, CUDA-core route
    ws = torch.empty((ws_bytes // 4,), device=x.device, dtype=torch.float32) if ws_bytes else None
    call("gram_fwd", KIND[kind], ptr(x), ptr(y), ptr(scale), scale.numel(), ptr(variance), ptr(out), ldo, n, n2, d,
         float(diag_add), ptr(diag_add_log), int(bool(lower_only)), ptr(ws), ws_bytes)
    return out


def gram_bwd_raw(kind, x, y, scale, variance, dk, need_x=True, need_y=True, need_scale=True, need_var=True,
                 symmetric=False):
    n, d = x.shape
    n2 = y.shape[0]
    dev = x.device
    dx = torch.empty_like(x) if need_x else None
    dy = torch.empty_like(y) if (need_y and not symmetric) else None
    dscale = torch.empty((scale.numel(),), device=dev, dtype=torch.float32) if need_scale else None
    dvar = torch.empty((1,), device=dev, dtype=torch.float32) if need_var else None
    dk = _f32c(dk)
    same = x.data_ptr() == y.data_ptr() and n == n2
    need_dy = (not symmetric) and (dy is not None or dscale is not None)
    ws_bytes = _lib.gram_workspace_bytes(n, n2, d, same, True, need_dy) if (dx is not None and (symmetric or dy is not None)) else 0
    ws = torch.empty((ws_bytes // 4,), device=dev, dtype=torch.float32) if ws_bytes else None
    call("gram_bwd", KIND[kind], ptr(x), ptr(y), ptr(scale), scale.numel(), ptr(variance), ptr(dk), dk.shape[1], n,
         n2, d, ptr(dx), ptr(dy), ptr(dscale), ptr(dvar), int(symmetric), ptr(ws), ws_bytes)
    return dx, dy, dscale, dvar


class _Gram(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, y, scale, variance, kind, diag_add):
        same = y is None
        x = _f32c(x)
        yy = x if same else _f32c(y)
        sc = _f32c(scale.reshape(-1))
        var = _f32c(variance.reshape(1))
        k = gram_raw(kind, x, yy, sc, var, diag_add=diag_add)
        ctx.save_for_backward(x, yy, sc, var)
        ctx.cfg = (kind, same, scale.shape, variance.shape)
        return k

    @staticmethod
    def backward(ctx, dk):
        x, yy, sc, var = ctx.saved_tensors
        kind, same, scale_shape, var_shape = ctx.cfg
        nx, ny, ns, nv = ctx.needs_input_grad[:4]
        if same:
            # K(x, x): the tensor feeds both the row and the column argument
            dx, dy, dscale, dvar = gram_bwd_raw(kind, x, yy, sc, var, dk, need_x=nx, need_y=nx, need_scale=ns,
                                                need_var=nv)
            if nx:
                dx = dx + dy
            dy = None
        else:
            dx, dy, dscale, dvar = gram_bwd_raw(kind, x, yy, sc, var, dk, need_x=nx, need_y=ny, need_scale=ns,
                                                need_var=nv)
        return (dx, dy, dscale.reshape(scale_shape) if dscale is not None else None,
                dvar.reshape(var_shape) if dvar is not None else None, None, None)


def gram(kind, x, y, scale, variance, diag_add=0.0):
    """Differentiable kernel matrix ``K(x, y) + diag_add * I`` (``y=None`` means ``K(x, x)``)."""
    return _Gram.apply(x, y, scale, variance, kind, float(diag_add))


# ------------------------------------------------------------------------------------------------
# kernel (4): Cholesky sweep and friends
# ------------------------------------------------------------------------------------------------
def potrf_ext_raw(e, n, n_ext_dense=0, ext_identity=False, want_diag=True, check=True):
    """In-place sweep over the stacked (n + r, ld) buffer ``e``.  Returns (diag(L), info tensor)."""
    dev = e.device
    ld = e.stride(0)
    ws_bytes = _lib.potrf_workspace_bytes(n, n_ext_dense)   # incl. the TMA scratch tail for large n
    ws = torch.empty((max(ws_bytes // 4, 1),), device=dev, dtype=torch.float32)
    info = torch.empty((1,), device=dev, dtype=torch.int32)
    diag = torch.empty((n,), device=dev, dtype=torch.float32) if want_diag else None
    call("potrf_ext", ptr(e), ld, n, n_ext_dense, int(ext_identity), ptr(diag), ptr(ws), ws_bytes, ptr(info))
    if check:
        check_info(info)
    return diag, info


def check_info(info):
    code = int(info.item())
    if code != 0:
        # same exception type as torch.cholesky on a non-PD input
        raise RuntimeError(
            "cholesky: the leading minor of order %d is not positive-definite (pinot_b200 potrf info=%d)"
            % (code, code)
        )


class _ExactGPLoss(torch.autograd.Function):
    """Fused exact-GP objective (pinot/regressors/gaussian_process_regressor.py:85-166).

    One Gram launch writes K + (eps + exp(log_sigma)) I straight into the sweep buffer; the sweep
    returns L, z = L^-1 y and L^-T; the gradient w.r.t. K is closed-form (no cholesky_backward).
    """

    @staticmethod
    def forward(ctx, x, y, scale, variance, log_sigma, kind, exact_logdet, check):
        x = _f32c(x)
        n, d = x.shape
        dev = x.device
        sc = _f32c(scale.reshape(-1))
        var = _f32c(variance.reshape(1))
        ls = _f32c(log_sigma.reshape(1))
        yv = _f32c(y.reshape(-1))
        assert yv.numel() == n, "exact GP: y must have one target per training point"
        need_grad = any(ctx.needs_input_grad[:5])
        ld = _pad4(n)
        rows = n + 1 + (n if need_grad else 0)
        e = torch.zeros((rows, ld), device=dev, dtype=torch.float32)
        gram_raw(kind, x, x, sc, var, out=e, ldo=ld, diag_add=EPSILON, diag_add_log=ls, lower_only=True)
        e[n, :n] = yv
        if need_grad:
            e[n + 1:, :n].diagonal().fill_(1.0)
        diag, info = potrf_ext_raw(e, n, n_ext_dense=1, ext_identity=need_grad, check=check)
        z = e[n, :n]
        out3 = torch.empty((3,), device=dev, dtype=torch.float32)
        call("gp_nll", ptr(diag), ptr(z), n, ptr(out3))
        ctx.save_for_backward(x, sc, var, ls, diag, e)
        ctx.cfg = (kind, int(exact_logdet), n, ld, scale.shape, variance.shape, log_sigma.shape, y.shape)
        ctx.info = info
        _ExactGPLoss.last_info = info
        nll = out3[1] if exact_logdet else out3[0]
        return nll.reshape(1, 1)

    @staticmethod
    def backward(ctx, g):
        x, sc, var, ls, diag, e = ctx.saved_tensors
        kind, exact_logdet, n, ld, scale_shape, var_shape, ls_shape, y_shape = ctx.cfg
        dev = x.device
        g = _f32c(g.reshape(1))
        z = e[n, :n]
        linv_t = e[n + 1:]
        alpha = torch.empty((n,), device=dev, dtype=torch.float32)
        call("gemv_rows", ptr(linv_t), ld, n, n, 1, ptr(z), ptr(alpha))
        dk = torch.empty((n, n), device=dev, dtype=torch.float32)
        ws_bytes = _lib.gp_nll_grad_workspace_bytes(n)
        ws = torch.empty((ws_bytes // 4,), device=dev, dtype=torch.float32) if ws_bytes else None
        call("gp_nll_grad", ptr(linv_t), ld, n, ptr(diag), ptr(alpha), ptr(g), exact_logdet, ptr(dk), n, ptr(ws),
             ws_bytes)
        del ws
        nx, ny, ns, nv, nl = ctx.needs_input_grad[:5]
        dx, _, dscale, dvar = gram_bwd_raw(kind, x, x, sc, var, dk, need_x=nx, need_y=False, need_scale=ns,
                                           need_var=nv, symmetric=True)
        if dx is not None:
            dx = dx * 2.0
        if dscale is not None:
            dscale = (dscale * 2.0).reshape(scale_shape)
        if dvar is not None:
            dvar = dvar.reshape(var_shape)
        dls = (torch.diagonal(dk).sum() * torch.exp(ls)).reshape(ls_shape) if nl else None
        dy = (alpha * g).reshape(y_shape) if ny else None
        return dx, dy, dscale, dvar, dls, None, None, None


def exact_gp_loss(x, y, scale, variance, log_sigma, kind="rbf", exact_logdet=False, check=True, return_info=False):
    """Exact-GP objective, shape (1, 1).  ``check=False`` leaves the factorisation status (LAPACK-style
    ``info``) on the device; with ``return_info`` it is returned for a later ``check_info``."""
    loss = _ExactGPLoss.apply(x, y, scale, variance, log_sigma, kind, bool(exact_logdet), bool(check))
    return (loss, _ExactGPLoss.last_info) if return_info else loss


@torch.no_grad()
def exact_gp_posterior(x_te, x_tr, y_tr, scale, variance, log_sigma, kind="rbf", full_cov=True, check=True):
    """Predictive mean and covariance (or marginal variance) of the exact GP,
    pinot/regressors/gaussian_process_regressor.py:168-233, from ONE sweep over
    ``[K_tr,tr + (eps + sigma^2) I ; y^T ; K_te,tr + eps I_rect]``.

    Returns (mean (n_te,), cov (n_te, n_te) or var (n_te,)).
    """
    x_te, x_tr = _f32c(x_te), _f32c(x_tr)
    n, d = x_tr.shape
    nte = x_te.shape[0]
    dev = x_tr.device
    sc = _f32c(scale.detach().reshape(-1))
    var = _f32c(variance.detach().reshape(1))
    ls = _f32c(log_sigma.detach().reshape(1))
    ld = _pad4(n)
    e = torch.zeros((n + 1 + nte, ld), device=dev, dtype=torch.float32)
    gram_raw(kind, x_tr, x_tr, sc, var, out=e, ldo=ld, diag_add=EPSILON, diag_add_log=ls, lower_only=True)
    e[n, :n] = _f32c(y_tr.detach().reshape(-1))
    gram_raw(kind, x_te, x_tr, sc, var, out=e[n + 1:], ldo=ld, diag_add=EPSILON)  # rectangular _perturb, :95
    potrf_ext_raw(e, n, n_ext_dense=1 + nte, ext_identity=False, want_diag=False, check=check)
    z = e[n, :n]
    vt = e[n + 1:]
    mean = torch.empty((nte,), device=dev, dtype=torch.float32)
    if not full_cov:
        v = torch.empty((nte,), device=dev, dtype=torch.float32)
        call("gp_predict_mean_var", ptr(vt), ld, nte, n, ptr(z), ptr(var), EPSILON, ptr(mean), ptr(v))
        return mean, v
    call("gp_predict_mean_var", ptr(vt), ld, nte, n, ptr(z), None, EPSILON, ptr(mean), None)
    cov = gram_raw(kind, x_te, x_te, sc, var, diag_add=EPSILON)
    # cov <- cov - Vt Vt^T (symmetric by construction, matching the 0.5 (C + C^T) of :221)
    call("sgemm", 0, 1, nte, nte, n, -1.0, ptr(vt), ld, ptr(vt), ld, 1.0, ptr(cov), nte)
    return mean, cov


class ExactGPPosterior:
    """Training-set factorisation kept for repeated prediction: one sweep over
    ``[K + (eps + sigma^2) I ; y^T]`` gives ``L`` and ``z = L^-1 y``; every later batch of test points then costs one
    Gram launch, one triangular solve ``V = K_te,tr L^-T`` against the cached factor (the same arithmetic as the
    reference's ``triangular_solve``, gaussian_process_regressor.py:215) and one fused row reduction (mean ``V z``,
    variance ``k** - |V_i|^2``) -- no second factorisation and, for the marginals, no n_te x n_te covariance.
    Equals the mean and (the diagonal of) the covariance of gaussian_process_regressor.py:168-233, rectangular
    ``_perturb`` included."""

    @torch.no_grad()
    def __init__(self, x_tr, y_tr, scale, variance, log_sigma, kind="rbf", check=True):
        self.kind = kind
        self.x_tr = _f32c(x_tr.detach())
        n = self.n = self.x_tr.shape[0]
        dev = self.x_tr.device
        self.scale = _f32c(scale.detach().reshape(-1))
        self.variance = _f32c(variance.detach().reshape(1))
        ls = _f32c(log_sigma.detach().reshape(1))
        self.ld = ld = _pad4(n)
        e = torch.zeros((n + 1, ld), device=dev, dtype=torch.float32)
        gram_raw(kind, self.x_tr, self.x_tr, self.scale, self.variance, out=e, ldo=ld, diag_add=EPSILON, diag_add_log=ls,
                 lower_only=True)
        e[n, :n] = _f32c(y_tr.detach().reshape(-1))
        potrf_ext_raw(e, n, n_ext_dense=1, ext_identity=False, want_diag=False, check=check)
        self.z = e[n, :n].clone()
        self.l = e                       # rows 0..n-1: L (lower triangle)

    def _solved_rows(self, x_te, row_offset=0):
        """``V = (K_te,tr + eps I_rect) L^-T`` for a batch of test points, (n_te, n)."""
        n = self.n
        g = gram_raw(self.kind, x_te, self.x_tr, self.scale, self.variance)
        if row_offset < n:   # _perturb of the rectangular K_te,tr: eps on its GLOBAL main diagonal (:95)
            g.diagonal(offset=int(row_offset)).add_(EPSILON)
        nbytes = _lib.trsm_rlt_workspace_bytes(n, g.shape[0])
        if nbytes and n % 4 == 0:
            # (one buffer per posterior, grown on demand: the split of L into it is repeated per call, 40 us at n = 4096)
            if getattr(self, "_trsm_ws", None) is None or self._trsm_ws.numel() * 4 < nbytes:
                self._trsm_ws = torch.empty((nbytes + 3) // 4, device=g.device, dtype=torch.float32)
            call("trsm_rlt_ws", ptr(self.l), self.ld, n, ptr(g), n, g.shape[0], ptr(self._trsm_ws), nbytes)
        else:
            call("trsm_rlt", ptr(self.l), self.ld, n, ptr(g), n, g.shape[0])
        return g

    @torch.no_grad()
    def mean_var(self, x_te, chunk=32768, row_offset=0):
        """Predictive mean and marginal variance, (n_te,) each, computed ``chunk`` test points at a time.
        ``row_offset``: index of ``x_te[0]`` within the full test set when the caller scores it in pieces
        (the rectangular ``_perturb`` sits on the GLOBAL main diagonal of K_te,tr)."""
        x_te = _f32c(x_te.detach())
        nte, n, dev = x_te.shape[0], self.n, x_te.device
        mean = torch.empty((nte,), device=dev, dtype=torch.float32)
        var = torch.empty((nte,), device=dev, dtype=torch.float32)
        for r0 in range(0, nte, chunk):
            r1 = min(nte, r0 + chunk)
            v = self._solved_rows(x_te[r0:r1], row_offset + r0)
            call("gp_predict_mean_var", ptr(v), n, r1 - r0, n, ptr(self.z), ptr(self.variance), EPSILON,
                 ptr(mean[r0:r1]), ptr(var[r0:r1]))
        return mean, var

    @torch.no_grad()
    def mean_cov(self, x_te):
        """Predictive mean (n_te,) and full covariance (n_te, n_te) from the cached factor:
        ``V = K_te,tr L^-T``, ``cov = K_te,te - V V^T`` (gaussian_process_regressor.py:212-221; symmetric by
        construction; its diagonal equals ``mean_var``'s variance bit for bit)."""
        x_te = _f32c(x_te.detach())
        nte, n, dev = x_te.shape[0], self.n, x_te.device
        v = self._solved_rows(x_te)
        mean = torch.empty((nte,), device=dev, dtype=torch.float32)
        var = torch.empty((nte,), device=dev, dtype=torch.float32)
        call("gp_predict_mean_var", ptr(v), n, nte, n, ptr(self.z), ptr(self.variance), EPSILON, ptr(mean), ptr(var))
        cov = gram_raw(self.kind, x_te, x_te, self.scale, self.variance, diag_add=EPSILON)
        call("sgemm", 0, 1, nte, nte, n, -1.0, ptr(v), n, ptr(v), n, 1.0, ptr(cov), nte)
        # The diagonal is a difference of nearly equal numbers (k** - |V_i|^2, often 10x smaller than either): it
        # takes the row reduction's value, summed in fp32 FMAs in one pass, instead of the tensor-core GEMM's
        # (3xTF32 products accumulated with truncation: ~1e-5 of |V_i|^2, i.e. ~1e-4 of a small variance).
        cov.diagonal().copy_(var)
        return mean, cov


class _CholSolve(torch.autograd.Function):
    """(K, B) -> (L, X = L^-1 B) with K = L L^T; differentiable (generic Cholesky backward built
    from L^-T, which the same sweep provides).  Used by the variational GP
    (pinot/regressors/gaussian_process_regressor.py:373-390, :466-473)."""

    @staticmethod
    def forward(ctx, k, b, check, info_sink=None):
        k = _f32c(k)
        n = k.shape[0]
        dev = k.device
        r = 0 if b is None else b.shape[1]
        need_grad = ctx.needs_input_grad[0] or (b is not None and ctx.needs_input_grad[1])
        ld = _pad4(n)
        e = torch.zeros((n + r + (n if need_grad else 0), ld), device=dev, dtype=torch.float32)
        e[:n, :n] = k
        if r:
            e[n:n + r, :n] = _f32c(b).t()
        if need_grad:
            e[n + r:, :n].diagonal().fill_(1.0)
        _, info = potrf_ext_raw(e, n, n_ext_dense=r, ext_identity=need_grad, want_diag=False, check=check)
        if info_sink is not None:
            info_sink.append(info)       # status stays on the device; the owner checks it when it synchronises
        l = e[:n, :n].clone()
        call("zero_upper", ptr(l), n, n)
        x = e[n:n + r, :n].t().contiguous() if r else None
        linv_t = e[n + r:, :n].contiguous() if need_grad else None
        ctx.save_for_backward(l, x, linv_t)
        ctx.has_b = b is not None
        if b is None:
            return l
        return l, x

    @staticmethod
    def backward(ctx, gl, gx=None):
        l, x, w = ctx.saved_tensors  # w = L^-T (upper triangular)
        n = l.shape[0]
        gl_tot = torch.zeros_like(l) if gl is None else torch.tril(_f32c(gl))
        gb = None
        if ctx.has_b and gx is not None:
            gb = matmul(w, _f32c(gx))                      # L^-T gX
            gl_tot = gl_tot - torch.tril(matmul(gb, x, False, True))
        gk = None
        if ctx.needs_input_grad[0]:
            p = torch.tril(matmul(l, gl_tot, True, False))  # Phi(L^T gL)
            p = p - 0.5 * torch.diag_embed(torch.diagonal(p))
            s = matmul(matmul(w, p), w, False, True)        # L^-T P L^-1
            gk = 0.5 * (s + s.t())
        if not ctx.has_b:
            return gk, None, None, None
        return gk, (gb if ctx.needs_input_grad[1] else None), None, None


def cholesky(k, check=True, info_sink=None):
    """Lower Cholesky factor (zeros above the diagonal), differentiable.  With ``check=False`` the
    LAPACK-style status stays on the device; ``info_sink`` (a list) receives that int tensor."""
    return _CholSolve.apply(k, None, check, info_sink)


def chol_solve(k, b, check=True, info_sink=None):
    """Returns ``(L, L^-1 b)`` for ``k = L L^T`` from one sweep, differentiable (``info_sink``: see
    :func:`cholesky`)."""
    return _CholSolve.apply(k, b, check, info_sink)


def expand_atom_features(z, fp, n_onehot=100, pad_pitch=False):
    """Dense ``(n_atoms, n_onehot + n_fp)`` fp32 atom features from their compact form (uint8 atomic number,
    fp16 or fp32 fingerprint columns): what pinot/graph.py:163-173 builds per atom on the host.

    ``pad_pitch``: lay the rows out with a pitch that is a multiple of 4 floats (117 -> 120, zeros in between) and
    return the ``(n, 117)`` view of that buffer.  Values and shape are the same; the 16-byte row alignment is what
    lets the first node update (``f_in``: Linear(117, H) + tanh) take the TMA-fed kernel, see :func:`node_update`."""
    assert z.is_cuda and fp.is_cuda and z.dtype == torch.uint8 and fp.dtype in (torch.float16, torch.float32)
    z, fp = z.contiguous(), fp.contiguous()
    n, n_fp = z.shape[0], (fp.shape[1] if fp.dim() == 2 else 0)
    nf = n_onehot + n_fp
    pitch = _pad4(nf) if pad_pitch else nf
    base = torch.empty((n, pitch), device=z.device, dtype=torch.float32)
    if pitch != nf:
        base[:, nf:].zero_()
    call("expand_atom_features", ptr(z), ptr(fp), fp.element_size(), n, n_onehot, n_fp, ptr(base), pitch)
    if pitch == nf:
        return base
    out = base[:, :nf]
    out._pb_pitch_base = base          # the zero-padded buffer this view lives in (node_update reads it)
    return out


def log_2pi():
    return math.log(2.0 * math.pi)
