mkdir -p gpurun_out
for w in 64 128 256 512 1024; do echo "outer block $w"; PINOT_B200_OUTER_BLOCK=$w timeout 300 python - <<'PY'
import os, sys, torch, time
sys.path.insert(0, os.getcwd())
from pinot_b200 import ops
DEV="cuda"
def t(fn, reps=3):
    fn(); torch.cuda.synchronize(); ts=[]
    for _ in range(reps):
        s,e=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True); s.record(); fn(); e.record(); torch.cuda.synchronize(); ts.append(s.elapsed_time(e))
    return sorted(ts)[len(ts)//2]
for n in (4200, 8192, 16384):
    x=0.1*torch.randn(n,128,device=DEV); sc=torch.zeros(128,device=DEV); var=torch.zeros((),device=DEV)
    k=ops.gram("rbf",x,None,sc,var,diag_add=0.05)
    ld=ops._pad4(n)
    e=torch.zeros((2*n+1,ld),device=DEV)
    def prep():
        e[:n,:n]=k; e[n+1:,:n].zero_(); e[n+1:,:n].diagonal().fill_(1.0)
    def run(ident):
        prep(); ops.potrf_ext_raw(e,n,n_ext_dense=1,ext_identity=ident,check=False)
    tp=t(prep)
    a=t(lambda: run(False))-tp; b=t(lambda: run(True))-tp
    run(True); l=torch.tril(e[:n,:n]); err=float(((l@l.t())-k).norm()/k.norm())
    print("n=%5d potrf %.2f ms  potrf+inverse %.2f ms  ||LL^T-K||/||K|| %.2e"%(n,a,b,err), flush=True)
PY
done
timeout 900 python -m pytest tests -m gpu -q --tb=line -p no:cacheprovider -k "potrf or chol or exact_gp or n16k or full_size" 2>&1 | tail -3
