"""ctypes binding of the C-ABI library ``libpinot_b200.so`` (declared in ``include/pinot_b200.h``).

There is no CPU fallback: every compute entry point requires CUDA tensors and the in-tree shared
library; a missing library or a CPU tensor raises immediately.
"""
import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
# PINOT_B200_LIB: another build of the same library (A/B timing of a kernel variant); the ABI check still applies
LIB_PATH = os.environ.get("PINOT_B200_LIB") or os.path.join(_HERE, "libpinot_b200.so")

_c_float_p = ctypes.c_void_p  # device pointers travel as raw addresses
_i64 = ctypes.c_int64
_i32 = ctypes.c_int
_f32 = ctypes.c_float
_ptr = ctypes.c_void_p

# name -> argtypes (restype is int unless listed in _RESTYPES); mirrors include/pinot_b200.h
SIGNATURES = {
    "pinot_b200_last_error": [],
    "pinot_b200_abi_version": [],
    "pinot_b200_launch_count": [],
    "pinot_b200_mp_csr_aggregate_fwd": [_ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _i64, _i64, _ptr],
    "pinot_b200_mp_csr_aggregate_bwd": [_ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _i64, _i64, _ptr],
    "pinot_b200_csr_build_workspace_bytes": [_i64, _i64],
    "pinot_b200_csr_build": [_ptr, _ptr, _i64, _i64, _ptr, _i64] + [_ptr] * 10 + [_ptr, _i64, _ptr],
    "pinot_b200_expand_atom_features": [_ptr, _ptr, _i32, _i64, _i64, _i64, _ptr, _i64, _ptr],
    "pinot_b200_gather_rows": [_ptr, _ptr, _ptr, _i64, _i64, _ptr],
    "pinot_b200_segment_readout_fwd": [_ptr, _ptr, _ptr, _i64, _i64, _i32, _ptr],
    "pinot_b200_segment_readout_bwd": [_ptr, _ptr, _ptr, _ptr, _i64, _i64, _i32, _ptr],
    "pinot_b200_node_update_fwd": [_ptr, _ptr, _ptr, _ptr, _i64, _i64, _i64, _i32, _i32, _i32, _ptr],
    "pinot_b200_node_update_bwd": [_ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _i64, _i64, _i64, _i32, _i32, _ptr],
    "pinot_b200_sgemm": [_i32, _i32, _i64, _i64, _i64, _f32, _ptr, _i64, _ptr, _i64, _f32, _ptr, _i64, _ptr],
    "pinot_b200_sgemm_nt_tma_workspace_bytes": [_i64, _i64, _i64],
    "pinot_b200_sgemm_nt_tma": [_i64, _i64, _i64, _f32, _ptr, _i64, _ptr, _i64, _f32, _ptr, _i64, _ptr, _i64, _ptr],
    "pinot_b200_tf32_split": [_ptr, _i64, _i64, _i64, _ptr, _i32, _ptr, _ptr, _ptr],
    "pinot_b200_sgemm_nt_presplit": [_i64, _i64, _i64, _f32, _ptr, _ptr, _ptr, _ptr, _f32, _ptr, _i64, _i32, _ptr, _ptr],
    "pinot_b200_gram_workspace_bytes": [_i64, _i64, _i64, _i32, _i32, _i32],
    "pinot_b200_gram_fwd": [_i32, _ptr, _ptr, _ptr, _i64, _ptr, _ptr, _i64, _i64, _i64, _i64, _f32, _ptr, _i32, _ptr, _i64, _ptr],
    "pinot_b200_gram_bwd": [_i32, _ptr, _ptr, _ptr, _i64, _ptr, _ptr, _i64, _i64, _i64, _i64, _ptr, _ptr, _ptr, _ptr, _i32, _ptr, _i64,
                            _ptr],
    "pinot_b200_potrf_workspace_bytes": [_i64],
    "pinot_b200_potrf_tma_scratch_bytes": [_i64, _i64],
    "pinot_b200_potrf_ext": [_ptr, _i64, _i64, _i64, _i32, _ptr, _ptr, _i64, _ptr, _ptr],
    "pinot_b200_potrf_ext_noalloc": [_ptr, _i64, _i64, _i64, _i32, _ptr, _ptr, _ptr],
    "pinot_b200_trsm_rlt": [_ptr, _i64, _i64, _ptr, _i64, _i64, _ptr],
    "pinot_b200_trsm_rlt_workspace_bytes": [_i64, _i64],
    "pinot_b200_trsm_rlt_ws": [_ptr, _i64, _i64, _ptr, _i64, _i64, _ptr, _i64, _ptr],
    "pinot_b200_zero_upper": [_ptr, _i64, _i64, _ptr],
    "pinot_b200_gemv_rows": [_ptr, _i64, _i64, _i64, _i32, _ptr, _ptr, _ptr],
    "pinot_b200_gp_nll": [_ptr, _ptr, _i64, _ptr, _ptr],
    "pinot_b200_gp_nll_grad_workspace_bytes": [_i64],
    "pinot_b200_gp_nll_grad": [_ptr, _i64, _i64, _ptr, _ptr, _ptr, _i32, _ptr, _i64, _ptr, _i64, _ptr],
    "pinot_b200_gp_predict_mean_var": [_ptr, _i64, _i64, _i64, _ptr, _ptr, _f32, _ptr, _ptr, _ptr],
}
_RESTYPES = {
    "pinot_b200_last_error": ctypes.c_char_p,
    "pinot_b200_potrf_workspace_bytes": ctypes.c_int64,
    "pinot_b200_gram_workspace_bytes": ctypes.c_int64,
    "pinot_b200_trsm_rlt_workspace_bytes": ctypes.c_int64,
    "pinot_b200_sgemm_nt_tma_workspace_bytes": ctypes.c_int64,
    "pinot_b200_potrf_tma_scratch_bytes": ctypes.c_int64,
    "pinot_b200_csr_build_workspace_bytes": ctypes.c_int64,
    "pinot_b200_gp_nll_grad_workspace_bytes": ctypes.c_int64,
    "pinot_b200_launch_count": ctypes.c_int64,
}

_lib = None
# number of C-ABI compute calls made so far (bench.py reports it as its kernel-launch claim basis)
call_count = 0


class PinotB200Error(RuntimeError):
    pass


def load():
    """Load the shared library (once) and declare every signature of the header."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise PinotB200Error(
            "pinot_b200: %s is missing. Build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "or `make -C pinot_b200/csrc`. There is no CPU fallback." % LIB_PATH
        )
    lib = ctypes.CDLL(LIB_PATH)
    for name, argtypes in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError here = header/library mismatch
        fn.argtypes = argtypes
        fn.restype = _RESTYPES.get(name, ctypes.c_int)
    _lib = lib
    return lib


def last_error():
    return load().pinot_b200_last_error().decode("utf-8", "replace")


def stream_ptr():
    return torch.cuda.current_stream().cuda_stream


def ptr(t):
    """Device address of a tensor (None -> NULL).  CPU tensors are rejected: no CPU path exists."""
    if t is None:
        return None
    if not t.is_cuda:
        raise PinotB200Error(
            "pinot_b200 kernels run on CUDA tensors only (got a %s tensor); there is no CPU fallback" % t.device
        )
    return t.data_ptr()


# optional timing probes: {entry point name: [(start_event, end_event), ...]} -- bench.py uses this to
# time chosen kernels with CUDA events on the launching stream inside the timed region
probe = None


def call(name, *args):
    """Invoke an int-returning entry point on the current stream; raise on a non-zero code."""
    global call_count
    lib = load()
    if probe is not None and name in probe:
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        rc = getattr(lib, "pinot_b200_" + name)(*args, stream_ptr())
        e.record()
        probe[name].append((s, e))
    else:
        rc = getattr(lib, "pinot_b200_" + name)(*args, stream_ptr())
    call_count += 1
    if rc != 0:
        raise PinotB200Error("pinot_b200_%s failed (code %d): %s" % (name, rc, last_error()))


def launch_count():
    """Kernels launched by the library so far in this process."""
    return int(load().pinot_b200_launch_count())


def potrf_workspace_bytes(n, n_ext_dense=None):
    """Minimum sweep workspace; with n_ext_dense given, including the TMA scratch tail (0 for small n)."""
    lib = load()
    base = int(lib.pinot_b200_potrf_workspace_bytes(int(n)))
    if n_ext_dense is None:
        return base
    return base + int(lib.pinot_b200_potrf_tma_scratch_bytes(int(n), int(n_ext_dense)))


def gp_nll_grad_workspace_bytes(n):
    return int(load().pinot_b200_gp_nll_grad_workspace_bytes(int(n)))


def trsm_rlt_workspace_bytes(n, r):
    """Scratch of the triangular solve with accurate wide updates (0: the plain schedule will run)."""
    return int(load().pinot_b200_trsm_rlt_workspace_bytes(int(n), int(r)))


def gram_workspace_bytes(n, n2, d, same, backward, need_dy=False):
    """Scratch of the tensor-core Gram route (0: the CUDA-core route will run)."""
    return int(load().pinot_b200_gram_workspace_bytes(int(n), int(n2), int(d), int(bool(same)), int(bool(backward)),
                                                      int(bool(need_dy))))
