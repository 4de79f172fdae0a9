"""Model registry with the reference's interface (rmvd/models/registry.py:7-53): ``@register_model`` on an
entrypoint function named after the model, ``list_models``, ``has_model``, ``get_model``."""

_entrypoints = {}
_trainable_entrypoints = {}


def register_model(arg=None, trainable=True):
    """Use as ``@register_model`` or ``@register_model(trainable=False)``."""
    def _register(entrypoint):
        _entrypoints[entrypoint.__name__] = entrypoint
        if trainable:
            _trainable_entrypoints[entrypoint.__name__] = entrypoint
        return entrypoint
    return _register(arg) if callable(arg) else _register


def list_models(trainable_only=False):
    return sorted(_trainable_entrypoints if trainable_only else _entrypoints)


def has_model(name, trainable_only=False):
    return name in (_trainable_entrypoints if trainable_only else _entrypoints)


def get_model(name):
    assert has_model(name), \
        f'The requested model "{name}" does not exist. Available models are: {" ".join(list_models())}'
    return _entrypoints[name]
