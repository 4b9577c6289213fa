"""CPU stand-ins for the C-ABI entry points of the GNN half (TEST INFRASTRUCTURE ONLY).

``-m "not gpu"`` tests that exercise host logic above the kernels -- the ``dgl`` facade under the
unmodified reference's training / active-learning loops -- cannot call CUDA kernels.  ``installed()``
swaps ``pinot_b200._lib.call`` / ``ptr`` for plain-torch functions with the same argument lists as the
entry points declared in include/pinot_b200.h (tensors travel in place of device pointers).  Nothing in
the product imports this file; the kernels themselves are checked against the oracle on the GPU.
"""
import contextlib

import torch


def _rows_of(indptr):
    counts = (indptr[1:] - indptr[:-1]).long()
    return torch.repeat_interleave(torch.arange(counts.numel()), counts)


def _aggregate(x, out, indptr, indices, gather_scale, row_scale, n, f):
    rows = _rows_of(indptr)
    src = indices.long()
    msg = x[src] if gather_scale is None else x[src] * gather_scale[src].reshape(-1, 1)
    acc = torch.zeros_like(out)
    acc.index_add_(0, rows, msg)
    out.copy_(acc if row_scale is None else acc * row_scale.reshape(-1, 1))


def mp_csr_aggregate_fwd(x, out, indptr, indices, src_scale, dst_scale, n, f):
    _aggregate(x, out, indptr, indices, src_scale, dst_scale, n, f)


def mp_csr_aggregate_bwd(dout, dx, indptr_t, indices_t, src_scale, dst_scale, n, f):
    # transposed graph: gather dout from the destinations (scaled by dst_scale), scale the row by src_scale
    _aggregate(dout, dx, indptr_t, indices_t, dst_scale, src_scale, n, f)


def segment_readout_fwd(x, out, seg_offsets, batch, f, mean):
    off = seg_offsets.long()
    for b in range(batch):
        seg = x[off[b]:off[b + 1]]
        out[b] = seg.mean(0) if (mean and seg.shape[0]) else seg.sum(0)


def segment_readout_bwd(dout, dx, seg_offsets, node_seg, n, f, mean):
    off = seg_offsets.long()
    g = dout[node_seg.long()]
    if mean:
        cnt = (off[1:] - off[:-1]).clamp(min=1).to(dout.dtype)
        g = g / cnt[node_seg.long()].reshape(-1, 1)
    dx.copy_(g)


def node_update_fwd(x, w, b, y, n, fin, fout, w_is_out_in, act, accumulate):
    z = x @ (w.t() if w_is_out_in else w)
    if b is not None:
        z = z + b
    if accumulate:
        z = z + y
    y.copy_(torch.tanh(z) if act else z)


def node_update_bwd(dy, y, x, w, dz, dx, dw, db, n, fin, fout, w_is_out_in, act):
    g = dy
    if act:
        dz.copy_(dy * (1.0 - y * y))
        g = dz
    if dx is not None:
        dx.copy_(g @ (w if w_is_out_in else w.t()))
    if dw is not None:
        dw.copy_(g.t() @ x if w_is_out_in else x.t() @ g)
    if db is not None:
        db.copy_(g.sum(0))


def gather_rows(v, index, out, n, f):
    out.copy_(v[index.long()])


TABLE = {f.__name__: f for f in (mp_csr_aggregate_fwd, mp_csr_aggregate_bwd, segment_readout_fwd, segment_readout_bwd,
                                 node_update_fwd, node_update_bwd, gather_rows)}


@contextlib.contextmanager
def installed():
    """Route ``_lib.call(name, ...)`` to the functions above while the block runs."""
    from pinot_b200 import _lib, ops
    saved = (_lib.call, _lib.ptr, ops.call, ops.ptr)

    def call(name, *args):
        if name not in TABLE:
            raise AssertionError("no CPU stand-in for pinot_b200_%s (GNN-half entry points only)" % name)
        with torch.no_grad():
            TABLE[name](*args)

    def ptr(t):
        return t

    _lib.call = ops.call = call
    _lib.ptr = ops.ptr = ptr
    try:
        yield
    finally:
        _lib.call, _lib.ptr, ops.call, ops.ptr = saved
