from . import base_kernel, rbf, matern  # noqa: F401
from .rbf import RBF  # noqa: F401
from .matern import Matern52  # noqa: F401
