"""robustmvd_b200: B200-native plane-sweep cost-volume path behind the rmvd model API.

``create_model``, ``prepare_custom_model``, ``list_models``, ``has_model``, ``register_model`` mirror ``rmvd``
(rmvd/__init__.py:18); the models' correlation / cost-volume / regression steps run in ``csrc/librmvd_b200.so``.
"""
__version__ = "0.1.0"

_API = ("create_model", "prepare_custom_model", "list_models", "has_model", "register_model")


def __getattr__(name):  # lazy: importing the package (e.g. for the synthetic inputs) must not need torch.nn models
    if name in _API:
        from . import models
        return getattr(models, name)
    raise AttributeError(name)
