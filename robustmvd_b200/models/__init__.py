"""Model API of the reference (rmvd/models/__init__.py): the two native models whose cost-volume path is the B200
hot path, plus the registry / factory functions."""
from .robust_mvd import robust_mvd, robust_mvd_5M, RobustMVD  # noqa: F401
from .mvsnet import mvsnet_train, MVSNet  # noqa: F401
from .factory import create_model, prepare_custom_model  # noqa: F401
from .registry import register_model, list_models, has_model  # noqa: F401
