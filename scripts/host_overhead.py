import os, sys, time, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pinot_b200 as pb
from pinot_b200 import synthetic
from pinot_b200.graph import BatchedGraph
dev = torch.device("cuda", 0)
torch.manual_seed(0)
net = pb.Net(pb.representation.Sequential(pb.representation.dgl_legacy.gn(), [128, "tanh"] * 3),
             output_regressor_class=pb.regressors.ExactGaussianProcessRegressor).to(dev)
with torch.no_grad():
    net.output_regressor.kernel.scale.fill_(-3.0)
g_host, y_host = synthetic.make_batch(4200, seed=2666)
g, y = g_host.to(dev), y_host.to(dev)
g.csr()
opt = torch.optim.Adam(net.parameters(), 1e-3)
def step():
    opt.zero_grad(set_to_none=True)
    loss = net.loss(g, y).sum(); loss.backward(); opt.step(); return loss
for _ in range(3): step()
torch.cuda.synchronize()
for name, fn in [("whole step", step)]:
    t0 = time.perf_counter(); l = fn(); t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
    print("%s: host enqueue %.2f ms, then GPU drain %.2f ms" % (name, (t1 - t0) * 1e3, (t2 - t1) * 1e3))
# pieces
def timed(label, fn, n=5):
    torch.cuda.synchronize(); ts = []
    for _ in range(n):
        t0 = time.perf_counter(); fn(); ts.append(time.perf_counter() - t0); torch.cuda.synchronize()
    print("%-28s host %.3f ms" % (label, 1e3 * sorted(ts)[n // 2]))
timed("forward (loss)", lambda: net.loss(g, y))
loss_holder = {}
def fb():
    opt.zero_grad(set_to_none=True); l = net.loss(g, y).sum(); l.backward()
timed("zero_grad+forward+backward", fb)
timed("opt.step", lambda: opt.step())
src_h, dst_h = (t.pin_memory() for t in g_host.edges()); feat_h = g_host.ndata["h"].pin_memory()
bnn, bne = g_host.batch_num_nodes().tolist(), g_host.batch_num_edges().tolist()
def upload():
    gs = BatchedGraph(src_h.to(dev, non_blocking=True), dst_h.to(dev, non_blocking=True), g_host.number_of_nodes(), bnn, bne, validate=False)
    gs.ndata["h"] = feat_h.to(dev, non_blocking=True)
    return gs
timed("upload (H2D submit + ctor)", upload)
gs = upload()
timed("csr build", lambda: (setattr(gs, "_csr", None), gs.csr()))
import cProfile, pstats
pr = cProfile.Profile(); pr.enable(); step(); pr.disable(); torch.cuda.synchronize()
pstats.Stats(pr).sort_stats("cumulative").print_stats(18)
