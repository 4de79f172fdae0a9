"""CPU oracle (numpy) of CVP-MVSNet's cost-volume functions -- TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Follows ``/root/reference/rmvd/models/blocks/cvp_mvsnet_components.py``: ``homo_warping`` :192-245 and ``proj_cost``
:375-457, including the in-place quirk of :392 (``volume_sq_sum = volume_sum.pow_(2)`` squares ``volume_sum`` too).
Pinned by golden vectors frozen from those functions run unmodified on CPU (``tools/make_cvp_golden.py``; ``.cuda()`` is
stubbed to the identity by the harness)."""
import numpy as np

from .planesweep import _f32, grid_unnormalize, F32


def _proj(ref_in, src_in, ref_ex, src_ex):
    last = np.array([[0, 0, 0, 1.0]], np.float32)
    out = []
    for b in range(len(src_in)):
        sp = np.concatenate([_f32(src_in[b]) @ _f32(src_ex[b])[:3], last], 0)
        rp = np.concatenate([_f32(ref_in[b]) @ _f32(ref_ex[b])[:3], last], 0)
        out.append((sp.astype(np.float64) @ np.linalg.inv(rp.astype(np.float64))).astype(np.float32))
    return np.stack(out)


def _warp(src, proj, depth):
    """src [B,C,H,W], proj [B,4,4], depth [B,D,H,W] -> [B,C,D,H,W] (bilinear, zeros, align_corners=False)."""
    src = _f32(src)
    B, C, H, W = src.shape
    D = depth.shape[1]
    x = np.arange(W, dtype=np.float32).reshape(1, 1, W)
    y = np.arange(H, dtype=np.float32).reshape(1, H, 1)
    out = np.zeros((B, C, D, H, W), np.float32)
    for b in range(B):
        R, T = proj[b, :3, :3], proj[b, :3, 3]
        rot = [(R[r, 0] * x + R[r, 1] * y + R[r, 2]).astype(np.float32) for r in range(3)]          # :414
        with np.errstate(all="ignore"):
            pz = [(rot[r] * depth[b] + T[r]).astype(np.float32) for r in range(3)]                  # :416-421  [D,H,W]
            xn = (pz[0] / pz[2]) / F32((W - 1) / 2) - F32(1)                                         # :422-424
            yn = (pz[1] / pz[2]) / F32((H - 1) / 2) - F32(1)
            ix, iy = grid_unnormalize(xn.astype(np.float32), W), grid_unnormalize(yn.astype(np.float32), H)
        x0, y0 = np.floor(ix), np.floor(iy)
        wx1, wy1 = ix - x0, iy - y0
        img = src[b].reshape(C, H * W)
        for dx, dy, wgt in ((0, 0, (1 - wy1) * (1 - wx1)), (1, 0, (1 - wy1) * wx1), (0, 1, wy1 * (1 - wx1)), (1, 1, wy1 * wx1)):
            xx, yy = x0 + dx, y0 + dy
            ok = (xx >= 0) & (xx <= W - 1) & (yy >= 0) & (yy <= H - 1)
            xi = np.clip(np.nan_to_num(xx), 0, W - 1).astype(np.int64)
            yi = np.clip(np.nan_to_num(yy), 0, H - 1).astype(np.int64)
            out[b] += img[:, yi * W + xi] * np.where(ok, wgt, 0).astype(np.float32)[None]
    return out


def homo_warping(src_feature, ref_in, src_in, ref_ex, src_ex, depth_hypos):
    """:192-245, ``depth_hypos [B,D]``."""
    B, _, H, W = src_feature.shape
    d = np.broadcast_to(_f32(depth_hypos)[:, :, None, None], (B, depth_hypos.shape[1], H, W))
    return _warp(src_feature, _proj(ref_in, src_in, ref_ex, src_ex), d)


def proj_cost(ref_feature, src_features, ref_in, src_in, ref_ex, src_ex, depth_hypos):
    """:375-457 with ``src_features`` the list of source maps at the level, ``src_in / src_ex [B,nsrc,...]``,
    ``depth_hypos [B,D,H,W]``."""
    ref = _f32(ref_feature)
    D = depth_hypos.shape[1]
    vol_sq = np.repeat(ref[:, :, None], D, 2) ** 2        # :391-392: volume_sum is squared in place, both start from ref^2
    vol_sum = vol_sq.copy()
    n = len(src_features)
    for v in range(n):
        w = _warp(src_features[v], _proj(ref_in, src_in[:, v], ref_ex, src_ex[:, v]), _f32(depth_hypos))
        vol_sum = vol_sum + w
        vol_sq = vol_sq + w ** 2
    return (vol_sq / F32(n + 1) - (vol_sum / F32(n + 1)) ** 2).astype(np.float32)   # :447-449
