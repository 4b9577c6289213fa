"""Host-side data helpers whose SEMANTICS the hot path depends on (pinot/data/utils.py)."""
from . import utils  # noqa: F401
from .packed import PackedBatch, PackedDataset, compact_features, expand_features_host  # noqa: F401,E402
