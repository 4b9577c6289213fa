"""configs[4] of BASELINE.json: batched graph message passing + Thompson-sampling acquisition over a
large synthetic pool, data-parallel over GPUs (one process per GPU, torchrun).

    python benchmarks/bo_bench.py [--pool 125000] [--train 4096] [--q 96] [--rounds 3]

Each rank holds `pool` molecules of the (pool * world)-molecule candidate set plus a replica of the
training set.  One acquisition round = embed the training set, factor the exact GP once, embed and
score this rank's pool (marginal posterior), draw q Thompson samples, global arg-max (one
all-gather), move the q winners' graphs out of the pool (device-side re-batching).  Prints one JSON
line with pool molecules scored per second (whole job) and the message-passing rate."""
import argparse
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import pinot_b200 as pb  # noqa: E402
from pinot_b200 import active, distributed as pbd, synthetic  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--pool", type=int, default=125000, help="pool molecules PER GPU (8 x 125k = 1M)")
    ap.add_argument("--train", type=int, default=4096)
    ap.add_argument("--q", type=int, default=96)
    ap.add_argument("--rounds", type=int, default=3)
    a = ap.parse_args()
    rank, world, local = pbd.init_from_env()
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    torch.manual_seed(0)
    net = pb.Net(pb.representation.Sequential(pb.representation.dgl_legacy.gn(), [128, "tanh", 128, "tanh", 128, "tanh"]),
                 output_regressor_class=pb.regressors.ExactGaussianProcessRegressor).to(dev)
    with torch.no_grad():
        net.output_regressor.kernel.scale.fill_(-3.0)
    pbd.broadcast_parameters(net)
    g_tr, y_tr = synthetic.make_batch(a.train, seed=11, device=dev)
    t0 = time.perf_counter()
    pool, _ = synthetic.make_batch(a.pool, seed=100 + rank, device=dev)
    gen_s = time.perf_counter() - t0
    pool.csr()
    net.g_last, net.y_last = g_tr, y_tr          # what Net.loss would have cached (pinot/net.py:191-193)
    offset = rank * a.pool

    def one_round():
        picks = active.thompson_sampling_sharded(net, pool, offset, q=a.q)
        mine = [int(p) - offset for p in picks.tolist() if offset <= int(p) < offset + a.pool]
        moved = pool.select(sorted(set(mine))) if mine else None
        return picks, moved

    one_round()
    torch.cuda.synchronize()
    if world > 1:
        torch.distributed.barrier()
    evs = []
    for _ in range(a.rounds):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        one_round()
        e.record()
        evs.append((s, e))
    torch.cuda.synchronize()
    ms = pbd.max_over_ranks(sum(s.elapsed_time(e) for s, e in evs) / a.rounds, dev)
    # message passing alone over the pool (3 GraphConv aggregations, H = 128)
    x = torch.randn(pool.number_of_nodes(), 128, device=dev)
    csr = pool.csr()
    pb.ops.aggregate(x, csr, "both")
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(3):
        pb.ops.aggregate(x, csr, "both")
    e.record()
    torch.cuda.synchronize()
    mp_ms = pbd.max_over_ranks(s.elapsed_time(e) / 3, dev)
    if rank == 0:
        print(json.dumps({
            "workload": "C5: Thompson-sampling acquisition over a %d-molecule pool (%d per GPU), training set %d, q=%d"
                        % (a.pool * world, a.pool, a.train, a.q),
            "n_gpus": world, "ms_per_round": ms, "pool_molecules_per_s": a.pool * world / (ms * 1e-3),
            "message_passing_ms_per_layer": mp_ms,
            "message_passing_molecules_per_s_per_layer": a.pool * world / (mp_ms * 1e-3),
            "pool_nodes_per_gpu": pool.number_of_nodes(), "pool_generation_s": gen_s, "scaling": "weak",
        }))
    if world > 1:
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
