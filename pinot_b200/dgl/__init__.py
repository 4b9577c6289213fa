"""A ``dgl``-shaped facade over pinot_b200's batched graph and message-passing kernels.

pinot imports ``dgl`` for exactly: ``DGLGraph`` (+ ``add_nodes`` / ``add_edges`` / ``ndata``,
pinot/graph.py:208-239), ``batch`` / ``unbatch`` (pinot/data/utils.py:260-276,
pinot/active/experiment.py:99-104), ``sum_nodes`` (pinot/representation/sequential.py:78) and
``dgl.nn.pytorch.conv.<Layer>`` (pinot/representation/dgl_legacy.py:40).  ``install_as_dgl()``
registers this package under those module names so the unmodified reference code runs on the
B200 kernels.
"""
import sys

from ..graph import BatchedGraph, batch, unbatch, sum_nodes, mean_nodes
from . import nn  # noqa: F401

DGLGraph = BatchedGraph
graph = BatchedGraph


class DGLError(Exception):
    pass


def install_as_dgl(force=False):
    """Make ``import dgl`` resolve to this facade (refuses to shadow a real DGL unless forced)."""
    here = sys.modules[__name__]
    existing = sys.modules.get("dgl")
    if existing is not None and existing is not here and not force:
        raise RuntimeError("a different 'dgl' module is already imported; pass force=True to replace it")
    sys.modules["dgl"] = here
    sys.modules["dgl.nn"] = nn
    sys.modules["dgl.nn.pytorch"] = nn.pytorch
    sys.modules["dgl.nn.pytorch.conv"] = nn.pytorch.conv
    return here
