"""Exact-GP marginal likelihood with the Gram matrix and its Cholesky sweep SHARDED over GPUs.

The reference has no multi-device code; this is SURVEY.md section 8(e)'s "Gram K / Cholesky / grad product"
row for the arithmetic of pinot/regressors/gaussian_process_regressor.py:85-166 (loss) and its backward.

Partitioning (one process per GPU, NCCL):

* the stacked sweep matrix  E = [K + (eps + sigma^2) I ; y^T ; I]  (see ops._ExactGPLoss) is cut into
  column blocks of width ``block`` (512); block J lives on rank J mod P, full height (only the rows that
  can ever be non-zero in it are stored: K rows J0.., the y row, identity rows 0..J1);
* right-looking sweep, one exchange per block: the owner factors its panel with the single-GPU sweep
  (``pinot_b200_potrf_ext`` on the panel: L_JJ, the solved rows below it, z_J and the rows of L^-T),
  broadcasts the solved panel, and every rank applies the rank-``block`` update to the blocks it owns
  (TMA-fed tensor-core GEMM from one hi/lo split of the panel);
* look-ahead: the owner of block J+1 updates and factors it FIRST and its broadcast travels while all
  ranks are still applying panel J to their other blocks (two CUDA streams per rank);
* backward: alpha = L^-T z and S = L^-T diag(L) L^-1 are sums over column blocks -> per-rank partial
  sums; S is reduced to the rank that owns each row block (the rank whose molecules those rows are), and
  the Gram backward runs on row blocks.  Hyper-parameter gradients come back as PER-RANK PARTIAL SUMS,
  to be summed by the same gradient all-reduce that data-parallel training already does.

Rows of x / y stay where the data-parallel GNN produced them: ``sharded_exact_gp_loss(x_local, y_local, ...)``
all-gathers the embeddings, and each rank gets d loss / d x_local back.
"""
import contextlib
import math

import torch
import torch.distributed as dist

from . import _lib, ops
from ._lib import call, ptr

EPSILON = ops.EPSILON
DEFAULT_BLOCK = 512


# --------------------------------------------------------------------------------------------------
# local kernels (C-ABI); tests on CPU substitute an object with the same methods
# --------------------------------------------------------------------------------------------------
class CudaKernels:
    """The per-rank arithmetic, all through the C-ABI library."""

    def gram_block(self, kind, x_rows, x_cols, sc, var, out, ld, ls):
        ops.gram_raw(kind, x_rows, x_cols, sc, var, out=out, ldo=ld, diag_add=EPSILON, diag_add_log=ls, lower_only=True)

    def factor_panel(self, blk, w, n_dense, identity):
        """In-place sweep of one panel; returns (diag(L_JJ), info tensor)."""
        return ops.potrf_ext_raw(blk, w, n_ext_dense=n_dense, ext_identity=identity, want_diag=True, check=False)

    def use_presplit(self, n):
        return n >= 4096 and _lib.gp_nll_grad_workspace_bytes(4096) > 0   # TMA path compiled in and enabled

    def split(self, x, rows, k, colscale, mode, hi, lo):
        call("tf32_split", ptr(x), x.stride(0), rows, k, ptr(colscale), mode, ptr(hi), ptr(lo))

    def gemm_presplit(self, m, n, k, alpha, a_hi, a_lo, b_hi, b_lo, beta, c, lower, counter):
        call("sgemm_nt_presplit", m, n, k, alpha, ptr(a_hi), ptr(a_lo), ptr(b_hi), ptr(b_lo), beta, ptr(c), c.stride(0),
             int(lower), ptr(counter))

    def gemm_nt(self, m, n, k, alpha, a, b, beta, c):
        call("sgemm", 0, 1, m, n, k, alpha, ptr(a), a.stride(0), ptr(b), b.stride(0), beta, ptr(c), c.stride(0))

    def nll_terms(self, diag, z, w, out3):
        call("gp_nll", ptr(diag), ptr(z), w, ptr(out3))

    def gemv(self, m, rows, cols, v, y):
        call("gemv_rows", ptr(m), m.stride(0), rows, cols, 0, ptr(v), ptr(y))

    def gram_bwd_rows(self, kind, x_rows, x_all, sc, var, dk_rows, need_scale, need_var):
        """Row-side pass only (gram_bwd mode 2): dx and dscale are HALF of what the rows contribute."""
        n, d = x_rows.shape
        dev = x_rows.device
        dx = torch.empty_like(x_rows)
        dscale = torch.empty((sc.numel(),), device=dev, dtype=torch.float32) if need_scale else None
        dvar = torch.empty((1,), device=dev, dtype=torch.float32) if need_var else None
        ws_bytes = _lib.gram_workspace_bytes(n, x_all.shape[0], d, False, True, False)
        ws = torch.empty((ws_bytes // 4,), device=dev, dtype=torch.float32) if ws_bytes else None
        call("gram_bwd", ops.KIND[kind], ptr(x_rows), ptr(x_all), ptr(sc), sc.numel(), ptr(var), ptr(dk_rows),
             dk_rows.stride(0), n, x_all.shape[0], d, ptr(dx), None, ptr(dscale), ptr(dvar), 2, ptr(ws), ws_bytes)
        return dx, dscale, dvar


class _Lanes:
    """Two CUDA streams per rank (panel lane: look-ahead update + factor + broadcast; update lane: the
    other rank-W updates) with the events that order them; a no-op on CPU tensors (gloo tests)."""

    def __init__(self, dev):
        self.cuda = dev.type == "cuda"
        if self.cuda:
            self.cur = torch.cuda.current_stream(dev)
            self.s = [torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)]
            for s in self.s:
                s.wait_stream(self.cur)

    def lane(self, i):
        return torch.cuda.stream(self.s[i]) if self.cuda else contextlib.nullcontext()

    def record(self, i):
        if not self.cuda:
            return None
        ev = torch.cuda.Event()
        ev.record(self.s[i])
        return ev

    def wait(self, i, ev):
        if self.cuda and ev is not None:
            self.s[i].wait_event(ev)

    def join(self):
        if self.cuda:
            for s in self.s:
                self.cur.wait_stream(s)


def _pad4(n):
    return (n + 3) // 4 * 4


class ShardedSweep:
    """State of one sharded factorisation (kept for the backward pass)."""

    def __init__(self, n, block, rank, world, group, dev, need_grad, kernels):
        self.n, self.W, self.rank, self.P, self.group, self.dev = n, block, rank, world, group, dev
        self.need_grad = need_grad
        self.k = kernels
        self.bounds = [(j0, min(j0 + block, n)) for j0 in range(0, n, block)]
        self.nJ = len(self.bounds)
        self.blk = {}      # J -> (rows_J, ld_J) buffer of an owned block
        self.diag = {}     # J -> diag(L_JJ)
        self.info = {}     # J -> device info of the panel sweep
        self.global_src = [dist.get_global_rank(group, r) if (group is not None and world > 1) else r
                           for r in range(world)]

    # ---- layout helpers -----------------------------------------------------------------------
    def owner(self, j):
        return j % self.P

    def owned(self, j):
        return j % self.P == self.rank

    def rows_of(self, j):
        j0, j1 = self.bounds[j]
        return (self.n - j0) + 1 + (j1 if self.need_grad else 0)

    def panel_rows(self, j):
        """Rows of the solved panel below the diagonal block: K rows j1.., y, identity rows 0..j1."""
        j0, j1 = self.bounds[j]
        return (self.n - j1) + 1 + (j1 if self.need_grad else 0)

    def z_row(self, j):
        return self.n - self.bounds[j][0]

    def winv_rows(self, j):
        """Rows of L^-T (global rows 0..j1) x the block's columns, after the sweep."""
        j0, j1 = self.bounds[j]
        r0 = (self.n - j0) + 1
        return self.blk[j][r0:r0 + j1]

    # ---- forward ------------------------------------------------------------------------------
    def assemble(self, kind, x, y, sc, var, ls):
        n = self.n
        for j, (j0, j1) in enumerate(self.bounds):
            if not self.owned(j):
                continue
            w = j1 - j0
            ld = _pad4(w)
            b = torch.zeros((self.rows_of(j), ld), device=self.dev, dtype=torch.float32)
            self.k.gram_block(kind, x[j0:], x[j0:j1], sc, var, b, ld, ls)
            b[n - j0, :w] = y[j0:j1]
            if self.need_grad:
                r0 = (n - j0) + 1 + j0
                b[r0:r0 + w, :w].diagonal().fill_(1.0)
            self.blk[j] = b

    def _factor(self, j):
        j0, j1 = self.bounds[j]
        w = j1 - j0
        n_dense = (self.n - j1) + 1 + (j0 if self.need_grad else 0)
        self.diag[j], self.info[j] = self.k.factor_panel(self.blk[j], w, n_dense, self.need_grad)

    def _update_block(self, j, jp, panel, split=None):
        """Block jp (owned) -= panel_j rows x (panel_j rows of jp's columns)^T."""
        j0, j1 = self.bounds[j]
        p0, p1 = self.bounds[jp]
        off = p0 - j1                                    # panel row of global K row p0
        wj, wp = j1 - j0, p1 - p0
        m = (self.n - p0) + 1 + (j1 if self.need_grad else 0)
        c = self.blk[jp]
        if split is None:
            self.k.gemm_nt(m, wp, wj, -1.0, panel[off:], panel[off:], 1.0, c)
        else:
            hi, lo, counter = split
            self.k.gemm_presplit(m, wp, wj, -1.0, hi[off:], lo[off:], hi[off:], lo[off:], 1.0, c, False, counter)

    def run(self):
        """The sweep proper.  Everything is enqueued asynchronously; the two lanes are joined at the end."""
        n, P = self.n, self.P
        lanes = _Lanes(self.dev)
        ldw = _pad4(self.W)
        max_rows = n + 1 if self.need_grad else n + 1
        recv = None
        if P > 1:
            recv = [torch.empty((max_rows * ldw,), device=self.dev, dtype=torch.float32) for _ in range(2)]
        presplit = self.k.use_presplit(n)
        if presplit:
            sp_hi = torch.empty((max_rows, ldw), device=self.dev, dtype=torch.float32)
            sp_lo = torch.empty((max_rows, ldw), device=self.dev, dtype=torch.float32)
            counter = torch.zeros((64,), device=self.dev, dtype=torch.int32)

        def panel_of(j):
            j0, j1 = self.bounds[j]
            w = j1 - j0
            if self.owned(j):
                return self.blk[j][w:]
            rows, ld = self.panel_rows(j), _pad4(w)
            return recv[j % 2][: rows * ld].view(rows, ld)

        def exchange(j):
            if P > 1:
                dist.broadcast(panel_of(j), src=self.global_src[self.owner(j)], group=self.group)

        panel_ready = [None] * self.nJ
        lane1_done = [None] * self.nJ
        with lanes.lane(0):
            if self.owned(0):
                self._factor(0)
            exchange(0)
            panel_ready[0] = lanes.record(0)
        for j in range(self.nJ):
            nxt = j + 1
            panel = panel_of(j)
            if nxt < self.nJ:
                with lanes.lane(0):
                    if j >= 1:
                        lanes.wait(0, lane1_done[j - 1])   # block nxt's earlier updates; receive buffer nxt % 2 is free
                    if self.owned(nxt):
                        self._update_block(j, nxt, panel)
                        self._factor(nxt)
                    exchange(nxt)
                    panel_ready[nxt] = lanes.record(0)
            with lanes.lane(1):
                lanes.wait(1, panel_ready[j])
                mine = [jp for jp in range(j + 2, self.nJ) if self.owned(jp)]
                if mine:
                    split = None
                    if presplit:
                        j0, j1 = self.bounds[j]
                        off = self.bounds[mine[0]][0] - j1
                        rows, wj = self.panel_rows(j), j1 - j0
                        ldk = _pad4(wj)
                        hi = sp_hi.view(-1)[: rows * ldk].view(rows, ldk)
                        lo = sp_lo.view(-1)[: rows * ldk].view(rows, ldk)
                        self.k.split(panel[off:], rows - off, wj, None, 0, hi[off:], lo[off:])
                        split = (hi, lo, counter)
                    for jp in mine:
                        self._update_block(j, jp, panel, split)
                lane1_done[j] = lanes.record(1)
        lanes.join()

    def first_failure(self):
        """Smallest failing global pivot (1-based, LAPACK style) over all ranks, 0 if none: device tensor."""
        big = 2 ** 62
        t = torch.full((1,), big, device=self.dev, dtype=torch.int64)
        for j, info in self.info.items():
            code = info.to(torch.int64)
            t = torch.minimum(t, torch.where(code > 0, code + self.bounds[j][0], torch.full_like(code, big)))
        if self.P > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MIN, group=self.group)
        return torch.where(t == big, torch.zeros_like(t), t)

    def nll_terms(self):
        """[0.5 z.z + trace(L) + 0.5 n log 2pi, 0.5 z.z + sum log L_ii + 0.5 n log 2pi, z.z], summed over ranks."""
        tot = torch.zeros((3,), device=self.dev, dtype=torch.float32)
        for j in self.blk:
            j0, j1 = self.bounds[j]
            out3 = torch.empty((3,), device=self.dev, dtype=torch.float32)
            self.k.nll_terms(self.diag[j], self.blk[j][self.z_row(j)], j1 - j0, out3)
            tot += out3
        if self.P > 1:
            dist.all_reduce(tot, group=self.group)
        return tot

    # ---- backward -----------------------------------------------------------------------------
    def alpha(self):
        """alpha = K^-1 y = L^-T z, summed over column blocks and ranks."""
        a = torch.zeros((self.n,), device=self.dev, dtype=torch.float32)
        for j in self.blk:
            j0, j1 = self.bounds[j]
            tmp = torch.empty((j1,), device=self.dev, dtype=torch.float32)
            self.k.gemv(self.winv_rows(j), j1, j1 - j0, self.blk[j][self.z_row(j)], tmp)
            a[:j1] += tmp
        if self.P > 1:
            dist.all_reduce(a, group=self.group)
        return a

    def grad_partial(self, exact_logdet):
        """This rank's part of 0.5 L^-T diag(L) L^-1 (or 0.5 K^-1): full symmetric (n, pad4(n)) matrix."""
        n = self.n
        n4 = _pad4(n)
        s = torch.zeros((n, n4), device=self.dev, dtype=torch.float32)
        presplit = self.k.use_presplit(n)
        if presplit:
            counter = torch.zeros((64,), device=self.dev, dtype=torch.int32)
        for j in self.blk:
            j0, j1 = self.bounds[j]
            w = j1 - j0
            wm = self.winv_rows(j)
            if presplit:
                ldk = _pad4(w)
                hi = torch.empty((j1, ldk), device=self.dev, dtype=torch.float32)
                lo = torch.empty((j1, ldk), device=self.dev, dtype=torch.float32)
                self.k.split(wm, j1, w, None if exact_logdet else self.diag[j], 0 if exact_logdet else 2, hi, lo)
                self.k.gemm_presplit(j1, j1, w, 0.5, hi, lo, hi, lo, 1.0, s, True, counter)
            else:
                ws = wm[:, :w] if exact_logdet else wm[:, :w] * self.diag[j]
                ws = ws.contiguous()
                self.k.gemm_nt(j1, j1, w, 0.5, ws, wm, 1.0, s)
        _mirror_lower(s, n)
        return s


def _mirror_lower(s, n, chunk=2048):
    """Copy the lower triangle of s[:, :n] onto the upper one, in place, by row chunks."""
    for a in range(0, n, chunk):
        b = min(a + chunk, n)
        d = s[a:b, a:b]
        low = torch.tril(d)
        d.copy_(low + torch.tril(low, -1).t())
        if b < n:
            s[a:b, b:n] = s[b:n, a:b].t()


def _gather_rows(t, group, world):
    """All-gather row blocks of possibly different lengths; returns (full tensor, list of sizes)."""
    if world == 1:
        return t, [t.shape[0]]
    n = torch.tensor([t.shape[0]], device=t.device, dtype=torch.long)
    sizes = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(sizes, n, group=group)
    sizes = [int(s.item()) for s in sizes]
    m = max(sizes)
    pad = torch.zeros((m,) + tuple(t.shape[1:]), device=t.device, dtype=t.dtype)
    pad[: t.shape[0]] = t
    out = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(out, pad, group=group)
    return torch.cat([o[:s] for o, s in zip(out, sizes)], dim=0), sizes


class _ShardedExactGPLoss(torch.autograd.Function):
    last_info = None

    @staticmethod
    def forward(ctx, x_local, y_local, scale, variance, log_sigma, kind, exact_logdet, group, block, kernels):
        world = dist.get_world_size(group) if (dist.is_available() and dist.is_initialized()) else 1
        rank = dist.get_rank(group) if world > 1 else 0
        xl = x_local.detach().to(torch.float32).contiguous()
        yl = y_local.detach().to(torch.float32).reshape(-1).contiguous()
        x, sizes = _gather_rows(xl, group, world)
        y, _ = _gather_rows(yl, group, world)
        n = x.shape[0]
        dev = x.device
        sc = scale.detach().to(torch.float32).reshape(-1).contiguous()
        var = variance.detach().to(torch.float32).reshape(1).contiguous()
        ls = log_sigma.detach().to(torch.float32).reshape(1).contiguous()
        need_grad = any(ctx.needs_input_grad[:5])
        sw = ShardedSweep(n, int(block), rank, world, group, dev, need_grad, kernels)
        sw.assemble(kind, x, y, sc, var, ls)
        sw.run()
        terms = sw.nll_terms()
        _ShardedExactGPLoss.last_info = sw.first_failure()
        ctx.sw = sw
        ctx.saved = (x, sc, var, ls, sizes)
        ctx.cfg = (kind, bool(exact_logdet), scale.shape, variance.shape, log_sigma.shape, y_local.shape)
        return (terms[1] if exact_logdet else terms[0]).reshape(1, 1)

    @staticmethod
    def backward(ctx, g):
        sw = ctx.sw
        x, sc, var, ls, sizes = ctx.saved
        kind, exact_logdet, scale_shape, var_shape, ls_shape, y_shape = ctx.cfg
        n, dev, P, rank = sw.n, sw.dev, sw.P, sw.rank
        g = g.detach().to(torch.float32).reshape(1)
        alpha = sw.alpha()
        s = sw.grad_partial(exact_logdet)
        r0 = sum(sizes[:rank])
        r1 = r0 + sizes[rank]
        if P > 1:
            # each row block is summed onto the rank whose molecules those rows are
            works, a = [], 0
            for r, sz in enumerate(sizes):
                if sz:
                    works.append(dist.reduce(s[a:a + sz], dst=sw.global_src[r], group=sw.group, async_op=True))
                a += sz
            for w in works:
                w.wait()
        dk = s[r0:r1]
        dk[:, :n].addr_(alpha[r0:r1], alpha, alpha=-0.5)
        dk.mul_(g)
        nx, ny, ns, nv, nl = ctx.needs_input_grad[:5]
        dx = dscale = dvar = dls = dy = None
        if r1 > r0:
            dxr, dscale, dvar = sw.k.gram_bwd_rows(kind, x[r0:r1].contiguous(), x, sc, var, dk, ns, nv)
            if nx:
                dx = dxr * 2.0                      # row side + (equal, by symmetry) column side
            if dscale is not None:
                dscale = (dscale * 2.0).reshape(scale_shape)
            if dvar is not None:
                dvar = dvar.reshape(var_shape)
            if nl:
                idx = torch.arange(r1 - r0, device=dev)
                dls = (dk[idx, idx + r0].sum() * torch.exp(ls)).reshape(ls_shape)
        else:
            dx = torch.zeros((0, x.shape[1]), device=dev) if nx else None
            dscale = torch.zeros(scale_shape, device=dev) if ns else None
            dvar = torch.zeros(var_shape, device=dev) if nv else None
            dls = torch.zeros(ls_shape, device=dev) if nl else None
        if ny:
            dy = (alpha[r0:r1] * g).reshape(y_shape)
        ctx.sw = None
        return dx, dy, dscale, dvar, dls, None, None, None, None, None


def sharded_exact_gp_loss(x_local, y_local, scale, variance, log_sigma, kind="rbf", exact_logdet=False, group=None,
                          block=DEFAULT_BLOCK, check=True, kernels=None):
    """Exact-GP objective over the UNION of all ranks' rows, shape (1, 1), identical on every rank.

    Gradients: d/dx_local and d/dy_local are this rank's rows of the full gradient; d/dscale, d/dvariance
    and d/dlog_sigma are per-rank partial sums -- sum them over ranks (GradientBucket.all_reduce does) to
    get the full gradient.  ``check`` raises torch's RuntimeError for a non-PD matrix on every rank."""
    kernels = kernels or CudaKernels()
    loss = _ShardedExactGPLoss.apply(x_local, y_local, scale, variance, log_sigma, kind, bool(exact_logdet), group,
                                     int(block), kernels)
    if check:
        code = int(_ShardedExactGPLoss.last_info.item())
        if code != 0:
            raise RuntimeError(
                "cholesky: the leading minor of order %d is not positive-definite (pinot_b200 sharded potrf info=%d)"
                % (code, code))
    return loss


def half_log_2pi_n(n):
    return 0.5 * n * math.log(2.0 * math.pi)
