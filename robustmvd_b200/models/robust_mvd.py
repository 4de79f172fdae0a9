"""Robust MVD model (rmvd/models/robust_mvd.py:26-158) with the plane-sweep correlation on the sm_100a library.

``forward(images, poses, intrinsics, keyview_idx)`` -> ``(pred, aux)``: encoder -> PlanesweepCorrelation (256 planes,
linear in inverse depth over (0.4 m, 1000 m), :71-80) -> learned fusion -> cost-volume encoder -> decoder;
``pred["depth"] = 1 / (invdepth + 1e-9)``, ``pred["depth_uncertainty"] = exp(log_b) / (invdepth + 1e-9)`` (:90-97).
Only ``corr_block`` is the hot path; the convolutions are stock cuDNN modules (``nets.py``).
"""
import math
import os

import numpy as np
import torch
import torch.nn as nn

from ..blocks.planesweep_corr import PlanesweepCorrelation
from ..utils import (ResizeInputs, exclude_index, get_torch_model_device, select_by_index, to_numpy, to_torch)
from .helpers import build_model_with_cfg
from .nets import DispnetContextEncoder, DispnetCostvolumeEncoder, DispnetDecoder, DispnetEncoder, LearnedFusion
from .registry import register_model

NUM_PLANES, MIN_DEPTH, MAX_DEPTH = 256, 0.4, 1000.0  # robust_mvd.py:77-79
WEIGHTS_5M = "https://lmb.informatik.uni-freiburg.de/people/schroepp/weights/robustmvd.pt"
WEIGHTS_600K = "https://lmb.informatik.uni-freiburg.de/people/schroepp/weights/robustmvd_600k.pt"


class RobustMVD(nn.Module):
    def __init__(self):
        super().__init__()
        self.encoder = DispnetEncoder()
        self.context_encoder = DispnetContextEncoder()
        self.corr_block = PlanesweepCorrelation()
        self.fusion_block = LearnedFusion()
        self.fusion_enc_block = DispnetCostvolumeEncoder()
        self.decoder = DispnetDecoder()
        self.init_weights()

    def init_weights(self):
        """Kaiming-normal (a = 0.2, leaky_relu) convolutions, zero biases, unit batch norms (robust_mvd.py:39-55)."""
        for m in self.modules():
            if isinstance(m, (nn.Conv2d, nn.ConvTranspose2d, nn.Conv3d, nn.ConvTranspose3d)):
                if m.weight is not None:
                    nn.init.kaiming_normal_(m.weight, a=0.2, nonlinearity="leaky_relu")
                if m.bias is not None:
                    nn.init.zeros_(m.bias)
            elif isinstance(m, (nn.BatchNorm2d, nn.BatchNorm3d)):
                if m.weight is not None:
                    nn.init.ones_(m.weight)
                if m.bias is not None:
                    nn.init.zeros_(m.bias)

    def forward(self, images, poses, intrinsics, keyview_idx, **_):
        image_key = select_by_index(images, keyview_idx)
        images_source = exclude_index(images, keyview_idx)
        intrinsics_key = select_by_index(intrinsics, keyview_idx)
        intrinsics_source = exclude_index(intrinsics, keyview_idx)
        source_to_key_transforms = exclude_index(poses, keyview_idx)
        fused_slope = None

        if image_key.is_cuda and len(images_source) > 0 and os.environ.get("RMVD_BATCH_ENCODER", "1") != "0":
            # key and source views through the encoder as ONE (1 + V) * N batch (robust_mvd.py:66-67 runs
            # it once per view): the same convolutions with 1 / (1 + V) of the launches; the per-view features are
            # contiguous slices of one tensor, which is also what the correlation call's single repack launch reads
            nb = image_key.shape[0]
            stack = torch.cat([image_key] + list(images_source), 0)
            if not torch.is_grad_enabled() and os.environ.get("RMVD_FUSED_ACT", "1") != "0":
                # second half of row f-2: the last encoder layer's LeakyReLU (dispnet_encoder.py:23-27) is applied by the
                # correlation call as it emits its bf16 operand layouts: the activated source maps are never written, the key
                # map's activated copy (context encoder, decoder skip connection) is written by the same kernel
                c1 = self.encoder.conv1(stack)
                c2 = self.encoder.conv2(c1)
                pre = self.encoder.conv3[0](c2)
                fused_slope = self.encoder.conv3[1].negative_slope
                enc_key = torch.empty_like(pre[:nb])
                all_enc_key = {"conv1": c1[:nb], "conv2": c2[:nb], "conv3a": enc_key}
                enc_pre_key, enc_sources = pre[:nb], list(pre[nb:].split(nb, 0))
            else:
                all_enc, enc = self.encoder(stack)
                all_enc_key = {k: v[:nb] for k, v in all_enc.items()}
                enc_key = enc[:nb]
                enc_sources = list(enc[nb:].split(nb, 0))
        else:
            all_enc_key, enc_key = self.encoder(image_key)
            enc_sources = [self.encoder(image_source)[1] for image_source in images_source]

        if fused_slope is not None:
            corrs, masks, _ = self.corr_block(
                feat_key=enc_pre_key, intrinsics_key=intrinsics_key, feat_sources=enc_sources,
                source_to_key_transforms=source_to_key_transforms, intrinsics_sources=intrinsics_source,
                num_sampling_points=NUM_PLANES, min_depth=MIN_DEPTH, max_depth=MAX_DEPTH,
                fused_leaky_relu=fused_slope, key_activated_out=enc_key)
        else:
            corrs, masks, _ = self.corr_block(
                feat_key=enc_key, intrinsics_key=intrinsics_key, feat_sources=enc_sources,
                source_to_key_transforms=source_to_key_transforms, intrinsics_sources=intrinsics_source,
                num_sampling_points=NUM_PLANES, min_depth=MIN_DEPTH, max_depth=MAX_DEPTH)
        ctx = self.context_encoder(enc_key)

        fused_corr, _ = self.fusion_block(corrs=corrs, masks=masks)
        all_enc_fused, enc_fused = self.fusion_enc_block(corr=fused_corr, ctx=ctx)
        dec = self.decoder(enc_fused=enc_fused, all_enc={**all_enc_key, **all_enc_fused})

        pred = {"depth": 1 / (dec["invdepth"] + 1e-9),
                "depth_uncertainty": torch.exp(dec["invdepth_log_b"]) / (dec["invdepth"] + 1e-9)}
        aux = dec
        aux["depth"] = pred["depth"]
        aux["depth_uncertainty"] = pred["depth_uncertainty"]
        return pred, aux

    def input_adapter(self, images, keyview_idx, poses, intrinsics, **_):
        """Resize to the next multiple of 64, ``image / 255 - 0.4``, intrinsics to relative units
        (robust_mvd.py:101-132)."""
        device = get_torch_model_device(self)
        orig_ht, orig_wd = images[0].shape[-2:]
        ht, wd = int(math.ceil(orig_ht / 64.0) * 64.0), int(math.ceil(orig_wd / 64.0) * 64.0)
        if (orig_ht, orig_wd) != (ht, wd):
            resized = ResizeInputs(size=(ht, wd))({"images": images, "intrinsics": intrinsics})
            images, intrinsics = resized["images"], resized["intrinsics"]
        scale_arr = np.array([[wd] * 3, [ht] * 3, [1.0] * 3], dtype=np.float32)
        intrinsics = [K / scale_arr for K in intrinsics]
        if (device.type == "cuda" and os.environ.get("RMVD_DEVICE_NORMALIZE", "1") != "0"
                and all(isinstance(im, np.ndarray) and im.dtype == np.float32 and im.shape == images[0].shape for im in images)):
            images = self._upload_and_normalize(images, device)
            keyview_idx, poses, intrinsics = to_torch((keyview_idx, poses, intrinsics), device=device)
        else:
            images = [image / 255.0 - 0.4 for image in images]
            images, keyview_idx, poses, intrinsics = to_torch((images, keyview_idx, poses, intrinsics), device=device)
        return {"images": images, "keyview_idx": keyview_idx, "poses": poses, "intrinsics": intrinsics}

    def _upload_and_normalize(self, images, device):
        """``image / 255.0 - 0.4`` (robust_mvd.py:118) on the device instead of in numpy: the raw float32 images go through ONE pinned
        staging buffer and one host-to-device copy, the two float32 operations (a true IEEE division by a device scalar -- torch
        turns a division by a Python scalar into a multiplication by the reciprocal -- and a subtraction) run as two launches on
        the whole stack.  Bit-identical to the numpy expression; ``RMVD_DEVICE_NORMALIZE=0`` restores the host path."""
        shape = (len(images),) + tuple(images[0].shape)
        st = getattr(self, "_stage", None)
        if st is None or tuple(st[0].shape) != shape or st[2].device != device:
            st = [torch.empty(shape, dtype=torch.float32).pin_memory(), None, torch.full((), 255.0, dtype=torch.float32, device=device)]
            self._stage = st
        if st[1] is not None:
            st[1].synchronize()                           # the previous call's copy out of the staging buffer has completed
        buf = st[0].numpy()
        for i, im in enumerate(images):
            np.copyto(buf[i], im)
        d = st[0].to(device, non_blocking=True)
        st[1] = torch.cuda.Event()
        st[1].record(torch.cuda.current_stream(device))
        d = torch.div(d, st[2])
        d -= 0.4
        return list(d.unbind(0))

    def output_adapter(self, model_output):
        pred, aux = model_output
        return to_numpy(pred), to_numpy(aux)


@register_model(trainable=False)
def robust_mvd_5M(pretrained=True, weights=None, train=False, num_gpus=1, **kwargs):
    weights = WEIGHTS_5M if (pretrained and weights is None) else None  # robust_mvd.py:144 (custom weights ignored)
    return build_model_with_cfg(model_cls=RobustMVD, weights=weights, train=train, num_gpus=num_gpus)


@register_model
def robust_mvd(pretrained=True, weights=None, train=False, num_gpus=1, **kwargs):
    weights = WEIGHTS_600K if (pretrained and weights is None) else weights
    return build_model_with_cfg(model_cls=RobustMVD, weights=weights, train=train, num_gpus=num_gpus)
