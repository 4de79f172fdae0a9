"""``create_model`` / ``prepare_custom_model`` with the reference's signatures (rmvd/models/factory.py:8-65)."""
import torch.nn as nn

from .helpers import add_run_function, place_model
from .registry import get_model


def create_model(name, pretrained=True, weights=None, train=False, num_gpus=1, **kwargs):
    """Create a registered model, attach ``run`` and ``name``.  ``weights`` (a path) overrides ``pretrained``."""
    model = get_model(name)(pretrained=pretrained, weights=weights, train=train, num_gpus=num_gpus, **kwargs)
    add_run_function(model)
    model.name = name
    return model


def prepare_custom_model(model, train=False, num_gpus=1):
    """A custom model must implement ``input_adapter``, ``forward`` and ``output_adapter``."""
    assert not isinstance(model, nn.DataParallel), \
        "Model must not be wrapped in nn.DataParallel before prepare_custom_model is called."
    model = place_model(model, train, num_gpus)
    add_run_function(model)
    return model
