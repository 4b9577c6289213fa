from . import kernels  # noqa: F401
from . import gaussian_process_regressor, neural_network_regressor  # noqa: F401
from .base_regressor import BaseRegressor  # noqa: F401
from .neural_network_regressor import NeuralNetworkRegressor  # noqa: F401
from .gaussian_process_regressor import (  # noqa: F401
    GaussianProcessRegressor,
    ExactGaussianProcessRegressor,
    ShardedExactGaussianProcessRegressor,
    VariationalGaussianProcessRegressor,
)
