"""``split`` and ``batch`` with the behaviour of pinot/data/utils.py:207-282, on ``pinot_b200.BatchedGraph``.

Kept quirks (they decide which molecules a training step sees):

* ``split`` sizes every part as ``int(len(ds) * share / sum(shares))`` -- truncation, so a few trailing
  entries can be left out of all parts (:221-224);
* ``batch`` counts BATCHES from the number of ENTRIES of ``ds`` (``len(ds) // batch_size``) but slices the
  flattened list of single graphs, and drops the remainder unless ``partial_batch`` (:248-278);
* ``shuffle`` reseeds Python's ``random`` with ``seed`` and shuffles ``ds`` IN PLACE (:253-255);
* targets come back stacked as ``(B,) + y.shape`` rows: ``(B, 1)`` for the usual one-target data sets.
"""
import random

import torch

from .. import graph as _graph


def split(ds, partition):
    """Consecutive slices of ``ds`` with sizes proportional to ``partition`` (integer truncation)."""
    n_data = len(ds)
    total = sum(partition)
    sizes = [int(n_data * share / total) for share in partition]
    parts, start = [], 0
    for size in sizes:
        parts.append(ds[start:start + size])
        start += size
    return parts


def batch(ds, batch_size, seed=2666, shuffle=False, partial_batch=False):
    """List of ``(BatchedGraph, y)`` with ``batch_size`` molecules each, in data-set order."""
    n_entries = len(ds)
    n_batches = n_entries // batch_size
    if shuffle:
        random.seed(seed)
        random.shuffle(ds)
    graphs, targets = [], []
    for g, y in ds:
        graphs.extend(_graph.unbatch(g))
        targets.extend(t.unsqueeze(0) for t in y)      # iterating a tensor yields its rows
    out = []
    for b in range(n_batches):
        lo, hi = b * batch_size, (b + 1) * batch_size
        out.append((_graph.batch(graphs[lo:hi]), torch.stack(targets[lo:hi], dim=0)))
    if partial_batch and n_entries % batch_size != 0:
        lo = n_batches * batch_size
        out.append((_graph.batch(graphs[lo:]), torch.stack(targets[lo:], dim=0)))
    return out
