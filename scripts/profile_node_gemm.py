"""One tall-and-skinny node GEMM (gemm_node.cu) per size, for an ncu capture:
ncu --set full --clock-control none -k regex:node_gemm -s 2 -c 2 python scripts/profile_node_gemm.py"""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pinot_b200._lib import call, ptr

torch.manual_seed(0)
for m in (105316, 1000000):
    a = torch.randn(m, 128, device="cuda")
    b = torch.randn(128, 128, device="cuda")
    c = torch.empty(m, 128, device="cuda")
    for _ in range(2):      # the first launch of each size is the warm-up (skipped by -s / picked by launch index)
        call("sgemm", 0, 0, m, 128, 128, 1.0, ptr(a), 128, ptr(b), 128, 0.0, ptr(c), 128)
    torch.cuda.synchronize()
