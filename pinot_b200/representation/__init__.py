from . import dgl_legacy, sequential  # noqa: F401
from .sequential import Sequential, SequentialMix  # noqa: F401
