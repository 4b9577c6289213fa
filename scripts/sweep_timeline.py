"""Development aid: per-task timeline of the single-launch sweep (library built with -DPB_SWEEP_TIMING).
    make -C pinot_b200/csrc clean && make -C pinot_b200/csrc EXTRA=-DPB_SWEEP_TIMING && python scripts/sweep_timeline.py 1128"""
import ctypes
import os
import sys
os.environ["PINOT_B200_SWEEP"] = "persistent"
import numpy as np
import torch
sys.path.insert(0, ".")
from pinot_b200 import _lib, ops
from pinot_b200._lib import call, ptr

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1128
ident = int(sys.argv[2]) if len(sys.argv) > 2 else 1
lib = _lib.load()
DEV = "cuda"
g = torch.Generator().manual_seed(n)
a = torch.randn(n, 128, generator=g).to(DEV)
k = a @ a.t() / 128 + torch.eye(n, device=DEV)
ld = ops._pad4(n)
rows = n + 1 + (n if ident else 0)
e0 = torch.zeros(rows, ld, device=DEV)
e0[:n, :n] = k
e0[n, :n] = 1.0
if ident:
    e0[n + 1:, :n].diagonal().fill_(1.0)
nbytes = _lib.potrf_workspace_bytes(n, 1)
ws = torch.empty(nbytes // 4, device=DEV)
info = torch.zeros(1, dtype=torch.int32, device=DEV)
wa = torch.randn(8192, 8192, device=DEV)
for _ in range(int(sys.argv[3]) if len(sys.argv) > 3 else 60):     # keep the GPU busy so that the clocks are up
    wa @ wa
for it in range(3):
    e = e0.clone()
    torch.cuda.synchronize()
    lib.pinot_b200_debug_sweep_log(None, 0, 1)
    call("potrf_ext", ptr(e), ld, n, 1, ident, None, ptr(ws), nbytes, ptr(info))
    torch.cuda.synchronize()
rec = np.dtype([("cta", "i4"), ("type", "i4"), ("I", "i4"), ("J", "i4"), ("t_pick", "u8"), ("t_start", "u8"), ("t_end", "u8"), ("ph", "u8", (8,)), ("clk_start", "u8"), ("clk_end", "u8")])
buf = np.zeros(1 << 16, dtype=rec)
fn = lib.pinot_b200_debug_sweep_log
fn.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int]
cnt = fn(buf.ctypes.data, len(buf), 0)
buf = buf[:cnt]
t0 = buf["t_pick"].min()
names = {1: "S0", 2: "U", 3: "S1", 4: "FAR"}
print("tasks", cnt, "span us", (buf["t_end"].max() - t0) / 1e3)
big = buf[(buf["t_end"] - buf["t_start"]) > 4000]
if len(big):
    mhz = (big["clk_end"] - big["clk_start"]).astype(float) / (big["t_end"] - big["t_start"]).astype(float) * 1e3
    print("SM clock during tasks (cycles / globaltimer ns): median %.0f MHz, min %.0f, max %.0f" % (np.median(mhz), mhz.min(), mhz.max()))
for ty in (1, 2, 3, 4):
    m = buf[buf["type"] == ty]
    if len(m):
        d = (m["t_end"] - m["t_start"]) / 1e3
        w = (m["t_start"] - m["t_pick"]) / 1e3
        diag = m[m["I"] == (m["J"] if ty != 4 else m["J"] // 1000)]
        dd = (diag["t_end"] - diag["t_start"]) / 1e3
        print("%-4s n=%5d exec us mean %.2f med %.2f max %.2f | wait-before mean %.2f | on diagonal tiles: n=%d mean %.2f"
              % (names[ty], len(m), d.mean(), np.median(d), d.max(), w.mean(), len(diag), dd.mean() if len(diag) else 0))
# chain view: the diagonal tiles' tasks and the sub-diagonal S1, in time order
print("chain (us from start): type I J cta pick start end")
sel = [r for r in buf if (r["type"] != 4 and r["I"] <= r["J"] + 1) or (r["type"] == 4 and r["I"] == r["J"] // 1000 and r["J"] % 1000 == r["J"] // 1000 - 1)]
sel.sort(key=lambda r: r["t_start"])
for r in sel[:60]:
    print("  %-4s I=%d J=%s cta=%d  %.1f %.1f %.1f" % (names[r["type"]], r["I"], r["J"], r["cta"], (r["t_pick"] - t0) / 1e3,
                                                       (r["t_start"] - t0) / 1e3, (r["t_end"] - t0) / 1e3))

print("phase stamps relative to task start (us), first diagonal-tile tasks of each type")
for ty in (1, 2, 3, 4):
    m = [r for r in buf if r["type"] == ty and r["I"] == (r["J"] if ty != 4 else r["J"] // 1000)][2:6]
    for r in m:
        print("  %-4s" % names[ty], " ".join("%6.2f" % ((int(x) - int(r["t_start"])) / 1e3) for x in r["ph"]), " end %.2f" % ((int(r["t_end"]) - int(r["t_start"])) / 1e3))
print("off-diagonal")
for ty in (1, 3, 4):
    m = [r for r in buf if r["type"] == ty and r["I"] != (r["J"] if ty != 4 else r["J"] // 1000)][10:13]
    for r in m:
        print("  %-4s" % names[ty], " ".join("%6.2f" % ((int(x) - int(r["t_start"])) / 1e3) for x in r["ph"]), " end %.2f" % ((int(r["t_end"]) - int(r["t_start"])) / 1e3))

fph = np.zeros(4, dtype=np.uint64)
fn(fph.ctypes.data, -1, 0)
print("factor64 cycles of thread 0 (column tile 2): (a) left-looking %d, (b) block factor + barrier %d, (c) replay + store + barrier %d" % tuple(int(v) for v in fph[:3]))

ws = np.zeros((4, 4, 6), dtype=np.int64)
fn(ws.ctypes.data, -2, 0)
print("factor64 per-warp stamps (cycles relative to warp 0's block start): loop top, after (a), after barrier 1, before barrier 2, after barrier 2")
for jb in range(4):
    base = ws[jb, 0, 0]
    for w in range(4):
        print("  jb=%d warp %d: " % (jb, w) + " ".join("%6d" % (int(v) - int(base)) for v in ws[jb, w, :5]))
