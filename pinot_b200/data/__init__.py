"""Host-side data helpers whose SEMANTICS the hot path depends on (pinot/data/utils.py)."""
from . import utils  # noqa: F401
