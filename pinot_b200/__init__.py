"""pinot_b200: a B200-native implementation of pinot's graph-convolution -> Gaussian-process path.

Public surface mirrors the reference (``pinot.Net``, ``pinot.representation``, ``pinot.regressors``);
arithmetic runs in the sm_100a kernels of ``libpinot_b200.so`` (``include/pinot_b200.h``).
"""
from . import graph  # noqa: F401
from . import dgl  # noqa: F401
from . import ops  # noqa: F401
from . import representation  # noqa: F401
from . import regressors  # noqa: F401
from . import net  # noqa: F401
from . import synthetic  # noqa: F401
from . import data  # noqa: F401
from . import active  # noqa: F401
from . import distributed  # noqa: F401
from . import captured  # noqa: F401
from .captured import CapturedTrainStep, CapturedBatchUpload  # noqa: F401
from .net import Net  # noqa: F401
from .graph import BatchedGraph, batch, unbatch, sum_nodes  # noqa: F401

__version__ = "0.1.0"
