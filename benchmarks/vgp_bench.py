"""configs[2] of BASELINE.json: 20k molecules, GraphSAGE 3x128 + VariationalGaussianProcessRegressor
with 512 inducing points; one optimisation step per minibatch (sizes 64 and 4096).

    python benchmarks/vgp_bench.py [--molecules 20000] [--inducing 512]
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import pinot_b200 as pb  # noqa: E402
from pinot_b200 import synthetic  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--molecules", type=int, default=20000)
    ap.add_argument("--inducing", type=int, default=512)
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    net = pb.Net(pb.representation.Sequential(pb.representation.dgl_legacy.gn("SAGEConv"),
                                              [128, "tanh", 128, "tanh", 128, "tanh"]),
                 output_regressor_class=pb.regressors.VariationalGaussianProcessRegressor,
                 n_inducing_points=a.inducing).to(dev)
    with torch.no_grad():
        net.output_regressor.kernel.scale.fill_(-3.0)
        net.output_regressor.x_tr.mul_(10.0)
    g, y = synthetic.make_batch(a.molecules, seed=2666, device=dev)
    opt = torch.optim.Adam(net.parameters(), 1e-3)
    for bs in (64, 4096):
        n_batches = a.molecules // bs
        batches = [(g.select(range(i * bs, (i + 1) * bs)), y[i * bs:(i + 1) * bs]) for i in range(min(n_batches, 6))]
        for gb, _ in batches:
            gb.csr()

        def step(gb, yb):
            opt.zero_grad(set_to_none=True)
            loss = net.loss(gb, yb).sum()
            loss.backward()
            opt.step()
            return loss

        for gb, yb in batches[:2]:
            step(gb, yb)
        torch.cuda.synchronize()
        evs = []
        for gb, yb in batches:
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            loss = step(gb, yb)
            e.record()
            evs.append((s, e))
        torch.cuda.synchronize()
        ms = sum(s.elapsed_time(e) for s, e in evs) / len(evs)
        line = {"workload": "C3: GraphSAGE 3x128 + VGP M=%d, minibatch %d of %d molecules" % (a.inducing, bs, a.molecules),
                "ms_per_step": ms, "molecules_per_s": bs / (ms * 1e-3), "loss": float(loss), "loss_dtype": str(loss.dtype)}
        # the same step replayed from a CUDA graph (minibatches padded to one shape could train this way): what the
        # device needs once the host's ~150 launches per step are out of the way
        try:
            import copy
            net_c = copy.deepcopy(net)
            opt_c = torch.optim.Adam(net_c.parameters(), 1e-3, capturable=True)
            cap = pb.CapturedTrainStep(net_c, opt_c, batches[0][0], batches[0][1])
            for _ in range(3):
                cap()
            torch.cuda.synchronize()
            evs = []
            for _ in range(10):
                s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                s.record()
                cap()
                e.record()
                evs.append((s, e))
            torch.cuda.synchronize()
            net_c.output_regressor.verify()
            line["captured_step_ms"] = sorted(s.elapsed_time(e) for s, e in evs)[5]
            line["captured_molecules_per_s"] = bs / (line["captured_step_ms"] * 1e-3)
            cap.release()
        except Exception as exc:  # noqa: BLE001 - report, keep the eager number
            line["captured_step_error"] = str(exc)[:200]
        print(json.dumps(line))


if __name__ == "__main__":
    main()
