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
        cur = torch.cuda.current_stream(dev)
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            for _ in range(max(int(warmup), 1)):
                self._eager()
            # undo the warm-up: parameters back, optimizer state zeroed in place (it must exist, at
            # fixed addresses, before the capture)
            with torch.no_grad():
                for p, s in zip(params, saved):
                    p.copy_(s)
                for st in optimizer.state.values():
                    for v in st.values():
                        if torch.is_tensor(v):
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
