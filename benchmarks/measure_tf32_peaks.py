"""TF32 tensor-pipe peaks on this GPU, measured the way the driver measured MEASURED_PEAKS.json
(8192^3 product, best of 10 = burst, back-to-back for 4 s = sustained; CUDA events):

  * torch.matmul fp32 with allow_tf32 (cuBLAS single-pass TF32)      -> tf32_dense_tflops[_sustained]
  * the library's own fp32-accurate 3xTF32 GEMMs (register-staged `sgemm` and TMA-fed `sgemm_nt_tma`)
    in fp32-equivalent TFLOP/s (2 N^3 per product; the tensor pipe executes 3x that)
  * torch.matmul fp32 without TF32 (FFMA) for reference

    python benchmarks/measure_tf32_peaks.py > gpurun_out/tf32_peaks.json
"""
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from pinot_b200 import _lib  # noqa: E402
from pinot_b200._lib import call, ptr  # noqa: E402

N = 8192
FLOP = 2.0 * N ** 3


def burst_and_sustained(fn, seconds=4.0):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    best = float("inf")
    for _ in range(10):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        best = min(best, s.elapsed_time(e))
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    k = 0
    s.record()
    while time.perf_counter() - t0 < seconds:
        for _ in range(8):
            fn()
        k += 8
        torch.cuda.synchronize()
    e.record()
    torch.cuda.synchronize()
    return FLOP / (best * 1e-3) / 1e12, FLOP * k / (s.elapsed_time(e) * 1e-3) / 1e12


def main():
    dev = "cuda"
    a = torch.randn(N, N, device=dev)
    b = torch.randn(N, N, device=dev)
    c = torch.empty(N, N, device=dev)
    out = {"gpu": torch.cuda.get_device_name(0), "n": N, "how": __doc__.split("\n\n")[0]}
    torch.backends.cuda.matmul.allow_tf32 = True
    out["tf32_dense_tflops"], out["tf32_dense_tflops_sustained"] = burst_and_sustained(lambda: torch.matmul(a, b, out=c))
    torch.backends.cuda.matmul.allow_tf32 = False
    out["fp32_ffma_tflops"], out["fp32_ffma_tflops_sustained"] = burst_and_sustained(lambda: torch.matmul(a, b, out=c), 2.0)
    out["own_3xtf32_sgemm_tflops"], out["own_3xtf32_sgemm_tflops_sustained"] = burst_and_sustained(
        lambda: call("sgemm", 0, 1, N, N, N, 1.0, ptr(a), N, ptr(b), N, 0.0, ptr(c), N))
    lib = _lib.load()
    ws_bytes = int(lib.pinot_b200_sgemm_nt_tma_workspace_bytes(N, N, N))
    if ws_bytes > 0:
        ws = torch.empty(ws_bytes // 4, device=dev)
        out["own_3xtf32_tma_tflops"], out["own_3xtf32_tma_tflops_sustained"] = burst_and_sustained(
            lambda: call("sgemm_nt_tma", N, N, N, 1.0, ptr(a), N, ptr(b), N, 0.0, ptr(c), N, ptr(ws), ws_bytes))
    out["note"] = ("own_* figures are fp32-equivalent (2 N^3 per product); the 3xTF32 split issues three TF32 MMAs per "
                   "product, so their tensor-pipe rate is 3x the figure and the ceiling is tf32_dense / 3")
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
