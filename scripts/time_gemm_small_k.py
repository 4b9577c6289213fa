import sys, json, torch
sys.path.insert(0, ".")
from pinot_b200._lib import call, ptr
DEV = "cuda"
def timed(fn, reps=8):
    fn(); torch.cuda.synchronize(); evs = []
    for _ in range(reps):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record(); evs.append((s, e))
    torch.cuda.synchronize()
    ts = sorted(a.elapsed_time(b) for a, b in evs); return ts[len(ts) // 2]
for M, N, K in ((65536, 2048, 128), (65536, 2048, 512), (32768, 3968, 128)):
    A = torch.randn(M, K, device=DEV); B = torch.randn(N, K, device=DEV); C = torch.randn(M, N, device=DEV)
    t_reg = timed(lambda: call("sgemm", 0, 1, M, N, K, -1.0, ptr(A), K, ptr(B), K, 1.0, ptr(C), N))
    ahi, alo = torch.empty_like(A), torch.empty_like(A); bhi, blo = torch.empty_like(B), torch.empty_like(B)
    cnt = torch.zeros(64, dtype=torch.int32, device=DEV)
    t_split = timed(lambda: (call("tf32_split", ptr(A), K, M, K, None, 0, ptr(ahi), ptr(alo)), call("tf32_split", ptr(B), K, N, K, None, 0, ptr(bhi), ptr(blo))))
    t_tma = timed(lambda: call("sgemm_nt_presplit", M, N, K, -1.0, ptr(ahi), ptr(alo), ptr(bhi), ptr(blo), 1.0, ptr(C), N, 0, ptr(cnt)))
    fl = 2.0 * M * N * K
    print(json.dumps({"M": M, "N": N, "K": K, "register_staged_ms": t_reg, "split_ms": t_split, "tma_presplit_ms": t_tma,
                      "tflops_register": fl / t_reg / 1e9, "tflops_tma": fl / t_tma / 1e9, "c_rmw_GBps_tma": 8.0 * M * N / t_tma / 1e6}))
