"""Train step captured in a CUDA graph.

pinot trains its exact-GP models full-batch: every optimisation step runs the same graphs through
`Net.loss` (pinot/app/experiment.py:78-89 with one batch; pinot/active/experiment.py:299-320).  The step
is then ~150 dependent kernel launches over fixed shapes, and the host's share (Python, ctypes, launch
calls) is what limits it once the kernels are fast.  ``CapturedTrainStep`` records
``zero_grad -> net.loss(g, y) -> backward -> optimizer.step()`` once and replays it with one call; new
data of the same shape is copied INTO the captured buffers with ``load``.
"""
import torch


class CapturedTrainStep:
    """``step = CapturedTrainStep(net, optimizer, g, y); loss = step()``.

    * ``optimizer`` must be graph-capturable (``torch.optim.Adam(..., capturable=True)``).
    * ``g`` / ``y`` are kept as the step's static inputs; ``load(g2, y2)`` overwrites them in place
      (same numbers of nodes, edges and graphs).
    * A non-positive-definite kernel matrix cannot raise from inside a replay: set the regressor's
      ``check_positive_definite = "deferred"`` (done here) and call ``verify()`` after reading the loss.
    * No autograd graph of an earlier eager step may still be alive (``del loss`` first): its
      AccumulateGrad nodes belong to another stream and invalidate the capture.
    * Construction runs ``warmup`` eager steps (to create optimizer state and cached graph views) and
      then restores parameters and optimizer state, so it does not train.
    """

    def __init__(self, net, optimizer, g, y, warmup=2, after_backward=None):
        dev = y.device
        assert dev.type == "cuda", "CapturedTrainStep needs CUDA tensors (there is no CPU path)"
        self.net, self.optimizer, self.g, self.y = net, optimizer, g, y
        self.after_backward = after_backward          # e.g. GradientBucket.allreduce for data-parallel runs
        reg = getattr(net, "output_regressor", None)
        if reg is not None and getattr(reg, "check_positive_definite", None) is True:
            reg.check_positive_definite = "deferred"
        g.csr()
        # Autograd graphs of earlier eager steps pin AccumulateGrad nodes to the stream they ran on and
        # invalidate a capture on another stream: drop the references this package holds (callers must not
        # keep an old loss tensor alive either).
        for obj in (reg, getattr(reg, "base_regressor", None)):
            for name in ("_x_tr", "_y_tr"):
                if obj is not None and hasattr(obj, name):
                    setattr(obj, name, None)
        params = [p for p in net.parameters()]
        saved = [p.detach().clone() for p in params]
        # optimizer state that exists BEFORE the first step (e.g. Adagrad's initial accumulator) is restored
        # value for value after the warm-up; state that the first step creates lazily starts from zeros
        # (Adam's moments and step counter, momentum buffers), which is what zeroing restores
        state0 = {p: {k: v.detach().clone() for k, v in st.items() if torch.is_tensor(v)}
                  for p, st in optimizer.state.items()}
        cur = torch.cuda.current_stream(dev)
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            for _ in range(max(int(warmup), 1)):
                self._eager()
            # undo the warm-up in place: parameters back, optimizer state back to its initial values (it
            # must exist, at fixed addresses, before the capture)
            with torch.no_grad():
                for p, s in zip(params, saved):
                    p.copy_(s)
                for p, st in optimizer.state.items():
                    for k, v in st.items():
                        if torch.is_tensor(v):
                            before = state0.get(p, {}).get(k)
                            if before is not None:
                                v.copy_(before)
                            else:
                                v.zero_()
        cur.wait_stream(side)
        torch.cuda.synchronize(dev)
        optimizer.zero_grad(set_to_none=True)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.loss = self._forward_backward()
            optimizer.step()

    def _forward_backward(self):
        loss = self.net.loss(self.g, self.y).sum()
        loss.backward()
        if self.after_backward is not None:
            self.after_backward()
        return loss

    def _eager(self):
        self.optimizer.zero_grad(set_to_none=True)
        loss = self._forward_backward()
        self.optimizer.step()
        return loss

    def load(self, g, y):
        """Copy a new batch of the same shape into the captured inputs (asynchronous, current stream)."""
        self.g.copy_from_(g)
        self.y.copy_(y, non_blocking=True)

    def __call__(self):
        self.graph.replay()
        return self.loss

    def release(self):
        """Destroy the captured graph and its memory pool.  Required before
        ``torch.distributed.destroy_process_group()`` when the step contains an NCCL collective: the
        communicator cannot be torn down while a graph that references it is alive."""
        torch.cuda.synchronize()
        if self.graph is not None:
            self.graph.reset()
        self.graph = None
        self.loss = None


class CapturedBatchUpload:
    """Host -> device transfer of one batch (edge list, node features, targets, per-graph node counts) plus the
    graph index build, captured once in a CUDA graph on a copy stream and replayed with one launch per step.

    The pinned HOST tensors given here are the source of every replay: a data loader refills them in place
    (same numbers of nodes / edges / graphs) and calls ``replay()``, which returns ``(graph, y, event)`` --
    device-resident, valid once ``event`` has completed, overwritten by the next replay of this object.
    Use two objects alternately (or copy out with ``CapturedTrainStep.load``) to overlap with compute.
    One driver call instead of ~25 per batch also takes the upload out of the way of a concurrently
    running compute graph."""

    def __init__(self, src_h, dst_h, feat_h, y_h, num_nodes, batch_num_nodes, batch_num_edges, stream, device=None):
        from .graph import BatchedGraph
        from . import ops
        # feat_h: the dense (n_atoms, F) fp32 matrix, or its compact form (z uint8, fp fp16, n_onehot) as kept by
        # pinot_b200.data.PackedBatch -- 13x fewer bytes over the link, expanded on the device inside the same graph
        compact = isinstance(feat_h, (tuple, list))
        feats = tuple(feat_h[:2]) if compact else (feat_h,)
        n_onehot = int(feat_h[2]) if compact and len(feat_h) > 2 else 100
        for t in (src_h, dst_h, y_h) + feats:
            assert t.is_pinned(), "CapturedBatchUpload needs pinned host tensors"
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.stream = stream
        self.bnn_h = torch.as_tensor(list(batch_num_nodes), dtype=torch.long).pin_memory()
        self.host = (src_h, dst_h, y_h, self.bnn_h) + feats
        self.dev = tuple(torch.empty_like(t, device=dev) for t in self.host)
        self.meta = (int(num_nodes), list(batch_num_nodes), list(batch_num_edges))
        self.h2d_bytes = sum(t.numel() * t.element_size() for t in self.host)

        def body():
            for d, h in zip(self.dev, self.host):
                d.copy_(h, non_blocking=True)
            g = BatchedGraph(self.dev[0], self.dev[1], self.meta[0], self.meta[1], self.meta[2], validate=False,
                             batch_num_nodes_device=self.dev[3])
            g.ndata["h"] = ops.expand_atom_features(self.dev[4], self.dev[5], n_onehot, pad_pitch=True) if compact else self.dev[4]
            g.csr()
            return g

        cur = torch.cuda.current_stream(dev)
        stream.wait_stream(cur)
        with torch.cuda.stream(stream):
            body()                                  # warm-up outside the capture (lazy initialisation, allocator)
        stream.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph, stream=stream):
            self.g = body()
        self.y = self.dev[2]

    def replay(self):
        with torch.cuda.stream(self.stream):
            self.graph.replay()
            ev = torch.cuda.Event()
            ev.record(self.stream)
        return self.g, self.y, ev

    def release(self):
        torch.cuda.synchronize()
        self.graph.reset()
        self.graph = None
