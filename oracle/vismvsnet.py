"""CPU oracle (numpy) of the Vis-MVSNet cost volume -- TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Follows ``/root/reference/rmvd/models/blocks/utils.py``: ``get_pixel_grids`` :155-160, ``homography_warping`` :179-189,
``interpolate`` :164-176, ``groupwise_correlation`` :71-89, and the call site
``rmvd/models/blocks/vis_mvsnet_singlestage.py:86-122, :230-242``.  Pinned by golden vectors frozen from those functions
run unmodified on CPU (``tools/make_vis_golden.py``; ``Tensor.cuda`` is stubbed to the identity there because
``get_pixel_grids`` hard-codes ``.cuda()``)."""
import numpy as np

from .planesweep import _f32, grid_unnormalize, F32


def homography_coords(Hs, h, w):
    """Unnormalised grid_sample coordinates ``ix, iy [n,d,h,w]`` of ``homography_warping`` for ``Hs [n,d,1|h,1|w,3,3]``."""
    Hs = _f32(Hs)
    x = (np.arange(w, dtype=np.float32) + F32(0.5)).reshape(1, 1, 1, w)
    y = (np.arange(h, dtype=np.float32) + F32(0.5)).reshape(1, 1, h, 1)
    with np.errstate(all="ignore"):
        q = [(Hs[..., r, 0] * x + Hs[..., r, 1] * y + Hs[..., r, 2]).astype(np.float32) for r in range(3)]   # :184
        den = q[2] + F32(1e-9)                                                                                # :185
        gx = ((q[0] / den) / F32(w) * F32(2) - F32(1)).astype(np.float32)                                     # :168-170
        gy = ((q[1] / den) / F32(h) * F32(2) - F32(1)).astype(np.float32)
        gx = np.clip(gx, F32(-1.1), F32(1.1))
        gy = np.clip(gy, F32(-1.1), F32(1.1))
        return grid_unnormalize(gx, w), grid_unnormalize(gy, h)


def homography_warp_volume(src, Hs):
    """``build_cost_volume`` (vis_mvsnet_singlestage.py:86-122): ``src [n,c,h,w]`` -> warped ``[n,c,d,h,w]``
    (bilinear, zeros padding, ``align_corners=False``)."""
    src = _f32(src)
    n, c, h, w = src.shape
    ix, iy = homography_coords(Hs, h, w)
    ix, iy = np.broadcast_to(ix, (n, Hs.shape[1], h, w)), np.broadcast_to(iy, (n, Hs.shape[1], h, w))
    x0, y0 = np.floor(ix), np.floor(iy)
    wx1, wy1 = ix - x0, iy - y0
    wx0, wy0 = F32(1) - wx1, F32(1) - wy1
    out = np.zeros((n, c) + ix.shape[1:], np.float32)
    for dx, dy, wgt in ((0, 0, wy0 * wx0), (1, 0, wy0 * wx1), (0, 1, wy1 * wx0), (1, 1, wy1 * wx1)):
        xx, yy = x0 + dx, y0 + dy
        ok = (xx >= 0) & (xx <= w - 1) & (yy >= 0) & (yy <= h - 1)
        xi = np.clip(xx, 0, w - 1).astype(np.int64)
        yi = np.clip(yy, 0, h - 1).astype(np.int64)
        for b in range(n):
            vals = src[b].reshape(c, h * w)[:, yi[b] * w + xi[b]]                      # [c,d,h,w]
            out[b] += vals * np.where(ok[b], wgt[b], F32(0))[None]
    return out


def groupwise_cost_volume(ref, src, Hs, groups=8):
    """``groupwise_correlation(ref_ncdhw, warped_src, groups, 1)`` (utils.py:71-89, vis_mvsnet_singlestage.py:242):
    ``[n,groups,d,h,w]``, SUM over the channels of a group."""
    ref = _f32(ref)
    warped = homography_warp_volume(src, Hs)
    n, c, d, h, w = warped.shape
    prod = ref[:, :, None] * warped
    return prod.reshape(n, groups, c // groups, d, h, w).sum(axis=2, dtype=np.float32)
