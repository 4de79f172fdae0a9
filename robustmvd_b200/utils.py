"""Host-side helpers of the model API (the thin part of ``rmvd/utils/utils.py`` and ``rmvd/data/transforms.py``
that the two models' adapters need).  Nothing here is on the hot path.

Reference behaviour followed: ``to_torch`` / ``to_numpy`` / ``numpy_collate`` (utils.py:126-262),
``select_by_index`` / ``exclude_index`` (:298-347), ``get_torch_model_device`` (:265-272), ``ResizeInputs`` and
``UpscaleInputsToNextMultipleOf`` (data/transforms.py:40-98), ``NormalizeImagesToMinMax`` /
``NormalizeImagesByShiftAndScale`` (:276-316), ``compute_depth_range`` (utils.py:22-41).
"""
import collections.abc
import math

import numpy as np
import torch

DEFAULT_DEPTH_RANGE = (0.1, 100.0)  # README.md:107, utils.py:22-41


def _is_np(x):
    return isinstance(x, (np.ndarray, np.generic))


def _rebuild(container, items):
    try:
        return type(container)(items)
    except TypeError:
        return list(items)


def to_torch(data, device=None):
    """Recursively convert numpy leaves to tensors on ``device``; tuples become lists (utils.py:126-168)."""
    if data is None:
        return None
    if isinstance(data, torch.Tensor):
        return data.to(device)
    if _is_np(data):
        if isinstance(data, np.ndarray) and data.dtype.kind in "SaUO":
            return data
        return torch.as_tensor(data, device=device)
    if isinstance(data, collections.abc.Mapping):
        return {k: to_torch(v, device) for k, v in data.items()}
    if isinstance(data, tuple) and hasattr(data, "_fields"):
        return type(data)(*(to_torch(v, device) for v in data))
    if isinstance(data, tuple):
        return [to_torch(v, device) for v in data]
    if isinstance(data, collections.abc.Sequence) and not isinstance(data, (str, bytes)):
        return _rebuild(data, [to_torch(v, device) for v in data])
    return data


def to_numpy(data):
    """Inverse of :func:`to_torch` (utils.py:228-262)."""
    if data is None:
        return None
    if isinstance(data, torch.Tensor):
        return data.detach().cpu().numpy()
    if _is_np(data):
        return data
    if isinstance(data, collections.abc.Mapping):
        return {k: to_numpy(v) for k, v in data.items()}
    if isinstance(data, tuple) and hasattr(data, "_fields"):
        return type(data)(*(to_numpy(v) for v in data))
    if isinstance(data, tuple):
        return [to_numpy(v) for v in data]
    if isinstance(data, collections.abc.Sequence) and not isinstance(data, (str, bytes)):
        return _rebuild(data, [to_numpy(v) for v in data])
    return data


def numpy_collate(batch):
    """Stack a list of samples into batched numpy structures (utils.py:171-225)."""
    if batch is None:
        return None
    elem = batch[0]
    if elem is None:
        assert all(e is None for e in batch)
        return None
    if isinstance(elem, torch.Tensor):
        return numpy_collate([b.detach().cpu().numpy() for b in batch])
    if isinstance(elem, np.ndarray):
        return elem[None] if len(batch) == 1 else np.stack(batch, 0)   # one sample: a view, not a copy
    if isinstance(elem, np.generic) or isinstance(elem, (float, int)):
        return np.array(batch)
    if isinstance(elem, (str, bytes)):
        return batch
    if isinstance(elem, collections.abc.Mapping):
        return {k: numpy_collate([d[k] for d in batch]) for k in elem}
    if isinstance(elem, tuple) and hasattr(elem, "_fields"):
        return type(elem)(*(numpy_collate(s) for s in zip(*batch)))
    if isinstance(elem, collections.abc.Sequence):
        n = len(elem)
        if not all(len(e) == n for e in batch):
            raise RuntimeError("each element in list of batch should be of equal size")
        cols = [numpy_collate(list(s)) for s in zip(*batch)]
        return cols if isinstance(elem, tuple) else _rebuild(elem, cols)
    raise TypeError(f"numpy_collate: unsupported element type {type(elem)}")


def _as_int_index(idx):
    if isinstance(idx, (int, np.integer)):
        return int(idx)
    if isinstance(idx, torch.Tensor) and idx.ndim == 0:
        return int(idx)
    return None


def _stack(items):
    return np.stack(items, 0) if isinstance(items[0], np.ndarray) else torch.stack(items, 0)


def select_by_index(views, idx):
    """``views[idx]``; ``idx`` may hold one index per batch sample (utils.py:298-320)."""
    i = _as_int_index(idx)
    if i is not None:
        return views[i]
    return _stack([views[int(i)][b] for b, i in enumerate(idx)])


def exclude_index(views, idx):
    """All views but ``idx``; ``idx`` may hold one index per batch sample (utils.py:323-347)."""
    i = _as_int_index(idx)
    if i is not None:
        return [v for k, v in enumerate(views) if k != i]
    per_sample = [[v[b] for k, v in enumerate(views) if k != int(i)] for b, i in enumerate(idx)]
    if per_sample and all(len(p) > 0 for p in per_sample):
        return [_stack(list(col)) for col in zip(*per_sample)]
    return per_sample


def get_torch_model_device(model):
    devices = {p.device for p in model.parameters()}
    if len(devices) != 1:
        raise RuntimeError("All model parameters need to be on the same device")
    return devices.pop()


def compute_depth_range(depth=None, invdepth=None, default_min_depth=DEFAULT_DEPTH_RANGE[0],
                        default_max_depth=DEFAULT_DEPTH_RANGE[1]):
    """utils.py:22-41: range of the valid values, else the project-wide default (0.1 m, 100 m)."""
    if depth is not None and (depth > 0).any():
        v = depth[depth > 0]
        return float(v.min()), float(v.max())
    if invdepth is not None and (invdepth > 0).any():
        v = invdepth[invdepth > 0]
        return 1.0 / float(v.max()), 1.0 / float(v.min())
    return default_min_depth, default_max_depth


# ---------------------------------------------------------------------------------------------------------
# input transforms used by the adapters (data/transforms.py:40-98, :276-316)
# ---------------------------------------------------------------------------------------------------------
def resize_image(image, ht, wd, order=1):
    """``skimage.transform.resize(image, [..., ht, wd], order)`` for up-scaling, restated with scipy: skimage >= 0.19
    implements it as ``ndimage.zoom(..., mode='mirror', grid_mode=True)`` on the float image (anti-aliasing only
    applies when down-scaling, which the adapters never do)."""
    from scipy import ndimage
    image = np.asarray(image)
    oh, ow = image.shape[-2:]
    if ht < oh or wd < ow:
        raise ValueError("resize_image only up-scales (the model adapters round sizes up)")
    zoom = [1.0] * (image.ndim - 2) + [ht / oh, wd / ow]
    out = ndimage.zoom(image.astype(np.float64), zoom, order=order, mode="mirror", grid_mode=True)
    return out.astype(image.dtype if image.dtype.kind == "f" else np.float64)


class ResizeInputs:
    """transforms.py:40-74: resize ``sample["images"]`` to (height, width) and scale ``sample["intrinsics"]`` rows."""

    def __init__(self, size, interpolation_order=1):
        self.height, self.width = size
        self.order = interpolation_order

    def __call__(self, sample):
        orig_ht, orig_wd = sample["images"][0].shape[-2:]
        ht, wd = self.height, self.width
        if sample.get("images") is not None:
            sample["images"] = [resize_image(im, ht, wd, self.order) for im in sample["images"]]
        if sample.get("intrinsics") is not None:
            scale = np.array([[wd / orig_wd] * 3, [ht / orig_ht] * 3, [1.0] * 3], dtype=np.float32)
            sample["intrinsics"] = [K * scale for K in sample["intrinsics"]]
        return sample


class UpscaleInputsToNextMultipleOf:
    """transforms.py:77-98."""

    def __init__(self, factor, interpolation_order=1):
        self.factor = factor
        self.order = interpolation_order

    def __call__(self, sample):
        orig_ht, orig_wd = sample["images"][0].shape[-2:]
        ht = int(math.ceil(orig_ht / self.factor) * self.factor)
        wd = int(math.ceil(orig_wd / self.factor) * self.factor)
        if (ht, wd) == (orig_ht, orig_wd):
            return sample
        return ResizeInputs((ht, wd), self.order)(sample)


def normalize_images_to_min_max(images, min_val=0.0, max_val=1.0):
    """transforms.py:276-290: [0,255] -> [min_val, max_val]."""
    return [(im / 255.0) * (max_val - min_val) + min_val for im in images]


def normalize_images_by_shift_and_scale(images, shift, scale):
    """transforms.py:293-316: per-channel ``(image - shift) / scale`` on channel-last views, float32 result."""
    shift = np.asarray(shift)
    scale = np.asarray(scale)
    out = []
    for im in images:
        cl = np.moveaxis(im, -3, -1)
        out.append(np.moveaxis((cl - shift) / scale, -1, -3).astype(np.float32))
    return out
