"""Per-step timeline of the launch-per-panel sweep (library built with EXTRA=-DPB_DEBUG_TIMING, selected through
PINOT_B200_LIB): %globaltimer stamps of tile (0,0) of every step's kernel and of the last tile-column-0 CTA to
finish its panel solve.  Prints the average over the middle steps of each interval of the chain."""
import ctypes
import os
import sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pinot_b200 import _lib, ops
from pinot_b200._lib import call, ptr

DEV = "cuda"
lib = _lib.load()
lib.pinot_b200_debug_steps.argtypes = [ctypes.c_void_p]
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4200
a = torch.randn(n, 128, device=DEV)
k = a @ a.t() / 128 + torch.eye(n, device=DEV)
ld = ops._pad4(n)
e0 = torch.zeros(2 * n + 1, ld, device=DEV)
e0[:n, :n] = k
e0[n, :n] = torch.randn(n, device=DEV)
e0[n + 1:, :n].diagonal().fill_(1.0)
nbytes = _lib.potrf_workspace_bytes(n, 1)
ws = torch.empty(nbytes // 4, device=DEV)
info = torch.zeros(1, dtype=torch.int32, device=DEV)
wa = torch.randn(8192, 8192, device=DEV)
for _ in range(30):
    wa @ wa
for _ in range(4):
    e = e0.clone()
    call("potrf_ext", ptr(e), ld, n, 1, 1, None, ptr(ws), nbytes, ptr(info))
torch.cuda.synchronize()
buf = (ctypes.c_ulonglong * (512 * 16))()
lib.pinot_b200_debug_steps(buf)
t = [[buf[i * 16 + j] for j in range(16)] for i in range(512)]
steps = [i for i in range(1, n // 64) if t[i][0] and t[i + 1][0]]
mid = steps[len(steps) // 4: 3 * len(steps) // 4]
names = {1: "set-up done (griddepcontrol.wait passed)", 2: "MMAs of tile (0,0) done", 3: "epilogue done", 4: "post-op begin",
         5: "factor done", 6: "flag published", 7: "tile (0,0) solve done", 15: "last column-0 CTA done"}


def avg(f):
    return sum(f(i) for i in mid) / len(mid)


print("n = %d, %d steps, averages over steps %d..%d, ns" % (n, len(steps), mid[0], mid[-1]))
print("   step period (start to next start)          %8.0f" % avg(lambda i: t[i + 1][0] - t[i][0]))
prev = 0
for j in (1, 2, 3, 4, 5, 6, 7, 15):
    print("   %-44s %8.0f  (+%.0f)" % (names[j], avg(lambda i: t[i][j] - t[i][0]), avg(lambda i: t[i][j] - t[i][prev])))
    prev = j
print("   inside the epilogue of tile (0,0), after the MMAs: first chunk out of TMEM %.0f, staged in shared memory %.0f, "
      "stored %.0f, second chunk stored %.0f" % tuple(avg(lambda i, j=j: t[i][j] - t[i][2]) for j in (11, 12, 13, 14)))
print("   next kernel's tile (0,0) start after that    %8.0f" % avg(lambda i: t[i + 1][0] - t[i][15]))
