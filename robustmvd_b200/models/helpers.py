"""``model.run`` and weight loading with the reference's behaviour (rmvd/models/helpers.py:21-172)."""
import collections.abc
from copy import deepcopy

import numpy as np
import torch

from ..utils import numpy_collate


def add_batch_dim(images, keyview_idx, poses=None, intrinsics=None, depth_range=None):
    return numpy_collate([(images, keyview_idx, poses, intrinsics, depth_range)])


def remove_batch_dim(batch):
    """Drop the leading batch dimension of every numpy leaf (helpers.py:28-62)."""
    if batch is None:
        return None
    if isinstance(batch, (np.ndarray, np.memmap)):
        return batch[0]
    if isinstance(batch, collections.abc.Mapping):
        return {k: remove_batch_dim(v) for k, v in batch.items()}
    if isinstance(batch, tuple) and hasattr(batch, "_fields"):
        return type(batch)(*(remove_batch_dim(x) for x in batch))
    if isinstance(batch, collections.abc.Sequence) and not isinstance(batch, (str, bytes)):
        items = [remove_batch_dim(x) for x in batch]
        try:
            return type(batch)(items)
        except TypeError:
            return items
    raise TypeError(f"remove_batch_dim: batch must contain numpy arrays, dicts or lists; found {type(batch)}")


def add_run_function(model):
    """Attach ``model.run(images, keyview_idx, poses, intrinsics, depth_range) -> (pred, aux)`` (helpers.py:65-89):
    un-batched inputs get a batch dimension that is removed again from the outputs."""

    @torch.no_grad()
    def run(images, keyview_idx, poses=None, intrinsics=None, depth_range=None, **_):
        no_batch_dim = images[0].ndim == 3
        if no_batch_dim:
            images, keyview_idx, poses, intrinsics, depth_range = add_batch_dim(images, keyview_idx, poses, intrinsics,
                                                                                depth_range)
        sample = model.input_adapter(images=images, keyview_idx=keyview_idx, poses=poses, intrinsics=intrinsics,
                                     depth_range=depth_range)
        pred, aux = model.output_adapter(model(**sample))
        if no_batch_dim:
            pred, aux = remove_batch_dim((pred, aux))
        return pred, aux

    model.run = run


def place_model(model, train=False, num_gpus=1):
    """train/eval mode and device placement (helpers.py:156-170, factory.py:43-62): 1 -> cuda, >1 -> DataParallel
    over ``range(num_gpus)`` with the adapters forwarded, <1 -> left on the host (only the convolutional parts can
    run there: the cost-volume blocks have no CPU path and raise)."""
    model.train() if train else model.eval()
    if num_gpus == 1:
        model = model.cuda()
    elif num_gpus > 1:
        parallel = torch.nn.DataParallel(model, device_ids=list(range(num_gpus))).cuda()
        parallel.input_adapter = model.input_adapter
        parallel.output_adapter = model.output_adapter
        model = parallel
    return model


def build_model_with_cfg(model_cls, cfg=None, weights=None, train=False, num_gpus=1, preprocess_weights_fct=None,
                         **kwargs):
    """Instantiate ``model_cls(**cfg, **kwargs)``, restore ``weights`` (file path or URL; the checkpoint's
    ``model_state_dict`` with any ``module.`` prefix stripped, ``strict=True``), set mode and device
    (helpers.py:104-172)."""
    if cfg is not None:
        cfg = deepcopy(cfg)
        cfg.update(kwargs)
        kwargs = cfg
    model = model_cls(**kwargs)
    if weights is not None:
        if weights.startswith("http"):
            state = torch.hub.load_state_dict_from_url(weights, map_location="cpu", progress=False, check_hash=False)
        else:
            state = torch.load(weights, map_location="cpu")
        if preprocess_weights_fct is not None:
            state = preprocess_weights_fct(state)
        state = {k.replace("module.", ""): v for k, v in state["model_state_dict"].items()}
        model.load_state_dict(state, strict=True)
    return place_model(model, train, num_gpus)
