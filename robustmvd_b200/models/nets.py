"""The convolutional networks either side of the hot path, as stock ``torch.nn`` modules (cuDNN): NOT part of the
B200 hot path, kept only so that ``create_model`` returns complete models and
reference checkpoints load with ``strict=True``.  Module / parameter names and construction order follow the reference
(so a seeded initialisation draws identical weights):

  DispnetEncoder              rmvd/models/blocks/dispnet_encoder.py:6-27
  DispnetContextEncoder       rmvd/models/blocks/dispnet_context_encoder.py:6-13
  LearnedFusion               rmvd/models/blocks/learned_fusion.py:5-54
  DispnetCostvolumeEncoder    rmvd/models/blocks/dispnet_costvolume_encoder.py:7-50
  DispnetDecoder              rmvd/models/blocks/dispnet_decoder.py:39-145
  FeatureNet, CostRegNet      rmvd/models/blocks/mvsnet_components.py:46-123
"""
import torch
import torch.nn as nn
import torch.nn.functional as F

LEAK = 0.2


def conv(cin, cout, kernel_size=3, stride=1):
    """Conv2d + LeakyReLU(0.2) (blocks/utils.py:14-28)."""
    return nn.Sequential(nn.Conv2d(cin, cout, kernel_size, stride, (kernel_size - 1) // 2, bias=True),
                         nn.LeakyReLU(LEAK, inplace=True))


class ReLUAndSigmoid(nn.Module):
    """Channel 0: ReLU; channels 1..: sigmoid stretched to [min, max] (blocks/utils.py:31-47)."""

    def __init__(self, inplace=False, min=0, max=1):
        super().__init__()
        self.inplace, self.min, self.max, self.range = inplace, min, max, max - min

    def forward(self, x):
        mean = F.relu(x[:, :1], inplace=self.inplace)
        rest = torch.sigmoid(x[:, 1:] * (4 / self.range)) * self.range + self.min
        return torch.cat([mean, rest], 1)


class DispnetEncoder(nn.Module):
    def __init__(self):
        super().__init__()
        self.conv1 = conv(3, 64, 7, 2)
        self.conv2 = conv(64, 128, 5, 2)
        self.conv3 = conv(128, 256, 3, 2)

    def forward(self, image):
        c1 = self.conv1(image)
        c2 = self.conv2(c1)
        c3 = self.conv3(c2)
        return {"conv1": c1, "conv2": c2, "conv3a": c3}, c3


class DispnetContextEncoder(nn.Module):
    def __init__(self):
        super().__init__()
        self.conv_redir = conv(256, 32, 1, 1)

    def forward(self, conv3):
        return self.conv_redir(conv3)


class LearnedFusion(nn.Module):
    """Per-view weight (3x3 conv 256->128, ReLU, 1x1 conv), softmax over views, mask-weighted average."""

    def __init__(self):
        super().__init__()
        self.view_weights, self.fused_corr, self.fused_mask = [], None, None
        self.corr_to_view_weight = nn.Sequential(nn.Conv2d(256, 128, 3, 1, 1, bias=True), nn.ReLU(inplace=True),
                                                 nn.Conv2d(128, 1, 1, 1, 0, bias=True))

    def reset(self):
        self.view_weights, self.fused_corr, self.fused_mask = [], None, None

    def forward(self, corrs, masks):
        self.reset()
        if len(corrs) == 1:
            self.fused_corr, self.fused_mask = corrs[0], masks[0]
            return self.fused_corr, self.fused_mask
        needs_grad = torch.is_grad_enabled() and (any(c.requires_grad for c in corrs) or
                                                  any(p.requires_grad for p in self.parameters()))
        if corrs[0].is_cuda and not needs_grad:   # training differentiates the torch expression below (forward-only kernel)
            return self._forward_cuda(corrs, masks)
        logits = torch.stack([self.corr_to_view_weight(c) for c in corrs], 0)  # V,N,1,h,w
        weights = F.softmax(logits, dim=0) + 1e-9
        self.view_weights = list(weights.unbind(0))
        masked = weights * torch.stack(masks, 0)                              # V,N,S,h,w
        wsum = masked.sum(0)
        self.fused_mask = (wsum != 0).float()
        csum = (torch.stack(corrs, 0) * masked).sum(0)
        self.fused_corr = csum / (wsum + 1e-9) * self.fused_mask
        return self.fused_corr, self.fused_mask


    @torch.no_grad()
    def _forward_cuda(self, corrs, masks):
        """learned_fusion.py:33-48 in one pass over corrs / masks (csrc/fusion.cu); the convolutions stay in cuDNN."""
        from .. import _lib
        V = len(corrs)
        N, S, h, w = corrs[0].shape
        f32c = lambda t: t if (t.dtype == torch.float32 and t.is_contiguous()) else t.float().contiguous()  # noqa: E731
        cs, ms = [f32c(c) for c in corrs], [f32c(m) for m in masks]
        logits = [f32c(self.corr_to_view_weight(c)) for c in cs]                # V x [N,1,h,w]
        dev = cs[0].device
        weights = torch.empty((V, N, 1, h, w), dtype=torch.float32, device=dev)
        fused_corr = torch.empty((N, S, h, w), dtype=torch.float32, device=dev)
        fused_mask = torch.empty_like(fused_corr)
        st = _lib.load().rmvd_view_fusion_fwd(_lib.ptr_array(logits), _lib.ptr_array(cs), _lib.ptr_array(ms), V, N, S, h, w,
                                              _lib.ptr(weights), _lib.ptr(fused_corr), _lib.ptr(fused_mask), _lib.stream_ptr())
        _lib.check(st, "rmvd_view_fusion_fwd")
        self.view_weights = list(weights.unbind(0))
        self.fused_corr, self.fused_mask = fused_corr, fused_mask
        return self.fused_corr, self.fused_mask


class DispnetCostvolumeEncoder(nn.Module):
    def __init__(self):
        super().__init__()
        self.conv3_1 = conv(256 + 32, 256)
        self.conv4 = conv(256, 512, stride=2)
        self.conv4_1 = conv(512, 512)
        self.conv5 = conv(512, 512, stride=2)
        self.conv5_1 = conv(512, 512)
        self.conv6 = conv(512, 1024, stride=2)
        self.conv6_1 = conv(1024, 1024)

    def forward(self, corr, ctx):
        out = {"merged": torch.cat([ctx, corr], 1)}
        x = out["merged"]
        for name in ("conv3_1", "conv4", "conv4_1", "conv5", "conv5_1", "conv6", "conv6_1"):
            x = getattr(self, name)(x)
            out[name] = x
        return out, x


def _iconv(cin, cout):
    return nn.Sequential(nn.Conv2d(cin + 2, cout, 3, 1, 1, bias=True), nn.LeakyReLU(LEAK, inplace=True))


def _pred(cin):
    return nn.Sequential(nn.Conv2d(cin, 2, 3, 1, 1, bias=True), ReLUAndSigmoid(inplace=True, min=-10, max=10))


def _deconv(cin, cout):
    return nn.Sequential(nn.ConvTranspose2d(cin, cout, 4, 2, 1, bias=True), nn.LeakyReLU(LEAK, inplace=True))


class DispnetDecoder(nn.Module):
    """Five refinement levels 1/64 -> 1/2: deconv, concat(skip, deconv, upsampled prediction), iconv, prediction of
    (inverse depth >= 0, log b in [-10, 10])."""

    SKIPS = ("conv5_1", "conv4_1", "conv3_1", "conv2", "conv1")

    def __init__(self):
        super().__init__()
        self.pred_0 = _pred(1024)
        cin = 1024
        for level, (cout, cskip) in enumerate(((512, 512), (256, 512), (128, 256), (64, 128), (32, 64)), start=1):
            setattr(self, f"deconv_{level}", _deconv(cin, cout))
            setattr(self, f"rfeat{level}", _iconv(cout + cskip, cout))
            setattr(self, f"pred_{level}", _pred(cout))
            cin = cout

    def forward(self, enc_fused, all_enc):
        preds = {}
        pred = self.pred_0(enc_fused)
        self.add_outputs(pred, preds)
        x = enc_fused
        for level, skip in enumerate(self.SKIPS, start=1):
            up = getattr(self, f"deconv_{level}")(x)
            pred_up = F.interpolate(pred, size=up.shape[-2:], mode="bilinear", align_corners=False).detach()
            x = getattr(self, f"rfeat{level}")(torch.cat((all_enc[skip], up, pred_up), 1))
            pred = getattr(self, f"pred_{level}")(x)
            self.add_outputs(pred, preds)
        return preds

    @staticmethod
    def add_outputs(pred, preds):
        mean, log_b = pred[:, 0:1], pred[:, 1:2]
        ent = torch.log(2 * torch.exp(log_b) + 1e-4) + 1
        preds.setdefault("invdepth_uncertainties_all", []).append(ent)
        preds.setdefault("invdepth_log_bs_all", []).append(log_b)
        preds.setdefault("invdepths_all", []).append(mean)
        preds["invdepth_uncertainty"], preds["invdepth_log_b"], preds["invdepth"] = ent, log_b, mean


# ---------------------------------------------------------------------------------------------------------
# MVSNet
# ---------------------------------------------------------------------------------------------------------
class ConvBnReLU(nn.Module):
    def __init__(self, cin, cout, kernel_size=3, stride=1, pad=1):
        super().__init__()
        self.conv = nn.Conv2d(cin, cout, kernel_size, stride=stride, padding=pad, bias=False)
        self.bn = nn.BatchNorm2d(cout)

    def forward(self, x):
        return F.relu(self.bn(self.conv(x)), inplace=True)


class ConvBnReLU3D(nn.Module):
    def __init__(self, cin, cout, kernel_size=3, stride=1, pad=1):
        super().__init__()
        self.conv = nn.Conv3d(cin, cout, kernel_size, stride=stride, padding=pad, bias=False)
        self.bn = nn.BatchNorm3d(cout)

    def forward(self, x):
        return F.relu(self.bn(self.conv(x)), inplace=True)


class FeatureNet(nn.Module):
    """8 convolutions, stride 4, 32 output channels."""

    def __init__(self):
        super().__init__()
        self.inplanes = 32
        spec = ((3, 8, 3, 1, 1), (8, 8, 3, 1, 1), (8, 16, 5, 2, 2), (16, 16, 3, 1, 1), (16, 16, 3, 1, 1),
                (16, 32, 5, 2, 2), (32, 32, 3, 1, 1))
        for i, s in enumerate(spec):
            setattr(self, f"conv{i}", ConvBnReLU(*s))
        self.feature = nn.Conv2d(32, 32, 3, 1, 1)

    def forward(self, x):
        for i in range(7):
            x = getattr(self, f"conv{i}")(x)
        return self.feature(x)


def _deconv3d(cin, cout):
    return nn.Sequential(nn.ConvTranspose3d(cin, cout, 3, padding=1, output_padding=1, stride=2, bias=False),
                         nn.BatchNorm3d(cout), nn.ReLU(inplace=True))


class CostRegNet(nn.Module):
    """3-D U-Net over the variance volume [B,32,D,h,w] -> [B,1,D,h,w]."""

    def __init__(self):
        super().__init__()
        self.conv0 = ConvBnReLU3D(32, 8)
        self.conv1 = ConvBnReLU3D(8, 16, stride=2)
        self.conv2 = ConvBnReLU3D(16, 16)
        self.conv3 = ConvBnReLU3D(16, 32, stride=2)
        self.conv4 = ConvBnReLU3D(32, 32)
        self.conv5 = ConvBnReLU3D(32, 64, stride=2)
        self.conv6 = ConvBnReLU3D(64, 64)
        self.conv7 = _deconv3d(64, 32)
        self.conv9 = _deconv3d(32, 16)
        self.conv11 = _deconv3d(16, 8)
        self.prob = nn.Conv3d(8, 1, 3, stride=1, padding=1)

    def forward(self, x):
        c0 = self.conv0(x)
        c2 = self.conv2(self.conv1(c0))
        c4 = self.conv4(self.conv3(c2))
        x = self.conv6(self.conv5(c4))
        x = c4 + self.conv7(x)
        x = c2 + self.conv9(x)
        x = c0 + self.conv11(x)
        return self.prob(x)
