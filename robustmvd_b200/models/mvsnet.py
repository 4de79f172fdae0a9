"""MVSNet model (rmvd/models/mvsnet.py:30-217) with the cost volume and the regression on the sm_100a library.

The reference materialises one warped volume per source view and ~12 full-volume elementwise passes (:123-136) and six
more passes for softmax / expectation / confidence (:140-158).  Here ``mvs_variance_volume`` writes the variance once
and ``mvs_regress`` reads the regularised cost once.  ``FeatureNet`` / ``CostRegNet`` stay stock cuDNN modules.

Deliberate difference: the reference overwrites the caller's ``poses`` with ``K @ pose`` in place (:82-83); this model
works on a copy (SURVEY.md 8b "Ownership").
"""
import torch
import torch.nn as nn

from ..blocks.utils import mvs_regress, mvs_variance_volume
from ..utils import (UpscaleInputsToNextMultipleOf, exclude_index, get_torch_model_device,
                     normalize_images_by_shift_and_scale, normalize_images_to_min_max, select_by_index, to_numpy,
                     to_torch)
from .helpers import build_model_with_cfg
from .nets import CostRegNet, FeatureNet
from .registry import register_model

DEFAULT_RANGE = (0.2, 100.0)  # mvsnet.py:52
IMAGENET_MEAN, IMAGENET_STD = [0.485, 0.456, 0.406], [0.229, 0.224, 0.225]


class MVSNet(nn.Module):
    def __init__(self, sample_in_inv_depth_space=False, num_sampling_steps=192):
        super().__init__()
        self.feature = FeatureNet()
        self.cost_regularization = CostRegNet()
        self.num_sampling_steps = num_sampling_steps
        self.sample_in_inv_depth_space = sample_in_inv_depth_space

    def depth_samples(self, depth_range, batch, device):
        """mvsnet.py:46-74: ``linspace(min, max, D)`` of the FIRST sample's range (default 0.2..100), evaluated on the
        host like the reference, repeated for the batch.  Inverse-depth sampling: the reference's ``[::-1]`` on a
        tensor raises; the intended ascending-depth schedule is produced with ``flip``."""
        lo, hi = DEFAULT_RANGE if depth_range is None else (float(depth_range[0][0]), float(depth_range[1][0]))
        D = self.num_sampling_steps
        if self.sample_in_inv_depth_space:
            d = 1 / torch.linspace(1 / hi, 1 / lo, D, dtype=torch.float32).flip(0)
        else:
            d = torch.linspace(lo, hi, D, dtype=torch.float32)
        return torch.stack(batch * [d]).to(device)

    @staticmethod
    def projection_matrices(poses, intrinsics, keyview_idx):
        """mvsnet.py:75-103: per view ``P = [K * diag(1/4, 1/4, 1) @ pose[:3,:4]; pose[3]]``, the key view inverted;
        returns ``[B, V, 4, 4]`` float32 with the key view first."""
        scale = None
        proj_mats = []
        for idx, (K_batch, pose_batch) in enumerate(zip(intrinsics, poses)):
            out = []
            for K, pose, key in zip(K_batch, pose_batch, keyview_idx):
                if scale is None:
                    scale = torch.tensor([[0.25] * 3, [0.25] * 3, [1.0] * 3], device=K.device)
                P = pose.clone()
                P[:3, :4] = torch.matmul(K * scale, P[:3, :4])
                if idx == int(key):
                    P = torch.inverse(P)
                out.append(P)
            proj_mats.append(torch.stack(out))
        ordered = [select_by_index(proj_mats, keyview_idx)] + exclude_index(proj_mats, keyview_idx)
        return torch.stack(ordered, 1).to(dtype=torch.float32)

    def forward(self, images, poses, intrinsics, keyview_idx, depth_range=None, **_):
        B = images[0].shape[0]
        device = images[0].device
        depth_samples = self.depth_samples(depth_range, B, device)            # B, D
        proj_mats = self.projection_matrices(poses, intrinsics, keyview_idx).to(device)
        views = [select_by_index(images, keyview_idx)] + exclude_index(images, keyview_idx)
        images = torch.stack(views, 1)                                        # B, V, 3, H, W
        V, H, W = images.shape[1], images.shape[3], images.shape[4]
        feats = self.feature(images.reshape(B * V, 3, H, W))
        feats = feats.reshape(B, V, *feats.shape[1:])                         # B, V, 32, h, w

        variance = mvs_variance_volume(feats, proj_mats, depth_samples)       # B, 32, D, h, w   (:123-136)
        cost_reg = self.cost_regularization(variance).squeeze(1)              # B, D, h, w
        reg = mvs_regress(cost_reg, depth_samples)                            # (:140-158)

        pred = {"depth": reg["depth"].unsqueeze(1), "depth_uncertainty": (1 - reg["confidence"]).unsqueeze(1)}
        return pred, {}

    def input_adapter(self, images, keyview_idx, poses=None, intrinsics=None, depth_range=None, masks=None):
        """Upscale to the next multiple of 32, [0,255] -> [0,1], ImageNet mean/std (mvsnet.py:170-199)."""
        device = get_torch_model_device(self)
        resized = UpscaleInputsToNextMultipleOf(32)({"images": images, "intrinsics": intrinsics})
        images = normalize_images_to_min_max(resized["images"], 0.0, 1.0)
        images = normalize_images_by_shift_and_scale(images, IMAGENET_MEAN, IMAGENET_STD)
        intrinsics = resized["intrinsics"]
        images, keyview_idx, intrinsics, poses, depth_range, masks = to_torch(
            (images, keyview_idx, intrinsics, poses, depth_range, masks), device=device)
        return {"images": images, "poses": poses, "intrinsics": intrinsics, "keyview_idx": keyview_idx,
                "depth_range": depth_range, "masks": masks}

    def output_adapter(self, model_output):
        pred, aux = model_output
        return to_numpy(pred), to_numpy(aux)


@register_model(trainable=False)
def mvsnet_train(pretrained=True, weights=None, train=False, num_gpus=1, **kwargs):
    assert not (pretrained and weights is None), "Pretrained weights are not available for this model."
    cfg = {"sample_in_inv_depth_space": False, "num_sampling_steps": 256}
    return build_model_with_cfg(model_cls=MVSNet, cfg=cfg, weights=None, train=train, num_gpus=num_gpus)
