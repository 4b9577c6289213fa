from . import conv  # noqa: F401
from .conv import GraphConv, SAGEConv  # noqa: F401
