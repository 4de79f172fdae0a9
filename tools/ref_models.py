"""Import the UNMODIFIED reference package ``rmvd`` from /root/reference on CPU (route B of SURVEY.md 8c).

Container-only tooling for ``tools/make_model_golden.py``.  The reference's optional dependencies that are absent here
(pytoml, skimage, matplotlib, kornia, easydict, torch._six) are replaced by minimal stand-ins BEFORE the import; none
of them touches the arithmetic of the two models under test (inputs are generated at sizes that need no resizing).
"""
import sys
import types

import numpy as np
import torch

REF_ROOT = "/root/reference"


def _module(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


def _create_meshgrid(height, width, normalized_coordinates=True, device=None, dtype=torch.float32):
    assert not normalized_coordinates
    ys, xs = torch.meshgrid(torch.linspace(0, height - 1, height, device=device, dtype=dtype),
                            torch.linspace(0, width - 1, width, device=device, dtype=dtype), indexing="ij")
    return torch.stack([xs, ys], -1).unsqueeze(0)


def _resize(image, output_shape, order=1, **_):
    from robustmvd_b200.utils import resize_image
    return resize_image(image, output_shape[-2], output_shape[-1], order)


class _Anything:
    def __init__(self, *a, **k):
        pass

    def __getattr__(self, name):
        return _Anything()

    def __call__(self, *a, **k):
        return _Anything()


def import_rmvd():
    if "rmvd" in sys.modules and getattr(sys.modules["rmvd"], "__file__", None):
        return sys.modules["rmvd"]
    for k in [k for k in sys.modules if k == "rmvd" or k.startswith("rmvd.") or k == "kornia" or k.startswith("kornia.")]:
        del sys.modules[k]
    _module("torch._six", string_classes=(str, bytes))
    try:
        import pytoml  # noqa: F401
    except ImportError:
        import tomllib
        _module("pytoml", load=lambda f: tomllib.loads(f.read()), loads=tomllib.loads)
    for name in ("skimage", "matplotlib", "kornia", "easydict"):
        try:
            __import__(name)
        except ImportError:
            pass
    if "skimage" not in sys.modules:
        sk = _module("skimage")
        sk.transform = _module("skimage.transform", resize=_resize)
    if "matplotlib" not in sys.modules:
        mpl = _module("matplotlib")
        for sub in ("pyplot", "colors", "cm"):
            setattr(mpl, sub, _module(f"matplotlib.{sub}", LinearSegmentedColormap=_Anything, ListedColormap=_Anything))
    if "kornia" not in sys.modules:
        k = _module("kornia")
        k.utils = _module("kornia.utils", create_meshgrid=_create_meshgrid)
    if "easydict" not in sys.modules:
        _module("easydict", EasyDict=dict)
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)
    import rmvd
    return rmvd


if __name__ == "__main__":
    r = import_rmvd()
    print(r.list_models())
