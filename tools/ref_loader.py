"""Load the UNMODIFIED reference hot-path files from /root/reference in isolation.

Used by ``tools/make_golden.py`` to freeze golden vectors under ``tests/golden/``, by ``tools/validate_oracle.py`` to pin
``oracle/`` against the reference code itself (both in the build container, where ``/root/reference`` exists), and by the
CPU baseline legs of ``bench.py`` (``cpu_baseline`` / ``--impl reference``), which on the GPU box find the same two files
under ``oracle/_ref/`` (``oracle/make_ref.py``).  Nothing in ``tests/`` or the product imports this module.

Route A of SURVEY.md section 8(c): stub ``rmvd.utils.to_torch`` and
``kornia.utils.create_meshgrid``, then ``importlib`` the two files as they lie.
"""
import importlib.util
import sys
import types

import numpy as np
import torch

import os

# The reference tree where it lies (build container), else the two hot-path files that oracle/make_ref.py copied, unmodified, into
# the git-ignored oracle/_ref/ (they travel to the GPU box with the snapshot; only bench.py's CPU baseline legs load them there).
_LOCAL = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref")
REF_ROOT = "/root/reference" if os.path.isdir("/root/reference/rmvd") else _LOCAL


def available():
    return os.path.exists(os.path.join(REF_ROOT, "rmvd", "models", "blocks", "planesweep_corr.py"))


def _to_torch(x, device=None):
    if isinstance(x, np.ndarray):
        x = torch.from_numpy(x)
    if isinstance(x, torch.Tensor) and device is not None:
        x = x.to(device)
    return x


def _create_meshgrid(height, width, normalized_coordinates=True, device=None, dtype=torch.float32):
    assert not normalized_coordinates
    xs = torch.linspace(0, width - 1, width, device=device, dtype=dtype)
    ys = torch.linspace(0, height - 1, height, device=device, dtype=dtype)
    yy, xx = torch.meshgrid(ys, xs, indexing="ij")
    return torch.stack([xx, yy], dim=-1).unsqueeze(0)  # 1,H,W,2


def _install_stubs():
    if "rmvd" not in sys.modules:
        rmvd = types.ModuleType("rmvd")
        utils = types.ModuleType("rmvd.utils")
        utils.to_torch = _to_torch
        rmvd.utils = utils
        sys.modules["rmvd"] = rmvd
        sys.modules["rmvd.utils"] = utils
    if "kornia" not in sys.modules:
        kornia = types.ModuleType("kornia")
        kutils = types.ModuleType("kornia.utils")
        kutils.create_meshgrid = _create_meshgrid
        kornia.utils = kutils
        sys.modules["kornia"] = kornia
        sys.modules["kornia.utils"] = kutils


def _load(name, rel):
    spec = importlib.util.spec_from_file_location(name, f"{REF_ROOT}/{rel}")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


_cache = {}


def planesweep_corr():
    _install_stubs()
    if "ps" not in _cache:
        _cache["ps"] = _load("_ref_planesweep_corr", "rmvd/models/blocks/planesweep_corr.py")
    return _cache["ps"]


def block_utils():
    _install_stubs()
    if "bu" not in _cache:
        _cache["bu"] = _load("_ref_block_utils", "rmvd/models/blocks/utils.py")
    return _cache["bu"]


def cvp_components():
    _install_stubs()
    if "cvp" not in _cache:
        _cache["cvp"] = _load("_ref_cvp_components", "rmvd/models/blocks/cvp_mvsnet_components.py")
    return _cache["cvp"]
