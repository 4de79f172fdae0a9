"""Oracle (test infrastructure): numpy restatement of the robust_mvd plane-sweep correlation.

Follows ``/root/reference/rmvd/models/blocks/planesweep_corr.py`` operation by operation.
All arithmetic is IEEE fp32 with every torch op rounded separately, exactly as the eager
reference does on CPU; the one fused operation is ATen's grid_sample unnormalisation
(see ``grid_unnormalize``).  Not imported by the product package.
"""
import numpy as np

F32 = np.float32
_ONE = F32(1.0)
_TWO = F32(2.0)


def _f32(x):
    return np.ascontiguousarray(np.asarray(x, dtype=np.float32))


def _fma32(a, b, c):
    """fp32 fused multiply-add emulated through fp64 (the product of two fp32 is exact in fp64)."""
    return (np.asarray(a, np.float64) * np.asarray(b, np.float64) + np.asarray(c, np.float64)).astype(np.float32)


# --------------------------------------------------------------------------------------
# a1: inverse-depth schedule
# --------------------------------------------------------------------------------------
def compute_sampling_invdepths(min_depth, max_depth, num_samples, sampling_type="linear_invdepth"):
    """planesweep_corr.py:524-555.  Returns ``[1 or N, S]`` float32, ascending inverse depth."""
    min_depth = _f32(min_depth).reshape(-1, 1)
    max_depth = _f32(max_depth).reshape(-1, 1)
    min_inv = _ONE / max_depth  # :539
    max_inv = _ONE / min_depth  # :540
    steps = np.arange(0, num_samples, dtype=np.float32)[None, :]  # :541
    if sampling_type == "linear_invdepth":
        # :546  min + (steps * (max - min)) / (S - 1): each op rounded
        out = min_inv + (steps * (max_inv - min_inv)) / F32(num_samples - 1)
    elif sampling_type == "linear_depth":
        # :550-553
        out = _ONE / (min_depth + (steps * (max_depth - min_depth)) / F32(num_samples - 1))
        out = out[:, ::-1]
    else:
        raise ValueError(sampling_type)
    return np.ascontiguousarray(out.astype(np.float32))


# --------------------------------------------------------------------------------------
# a3: epipolar coefficients
# --------------------------------------------------------------------------------------
def epipolar_scalars(intrinsics_key, intrinsics_source, source_to_key_transform, height, width,
                     height_source, width_source):
    """The twelve per-sample scalars a..m of planesweep_corr.py:239-275 as a dict of ``[N]`` arrays.

    Intrinsics are RELATIVE (fx/W ...), transforms are 4x4 key->source.
    """
    Kk = _f32(intrinsics_key)
    Ks = _f32(intrinsics_source)
    T = _f32(source_to_key_transform)
    fx = Kk[:, 0, 0] * F32(width)
    fy = Kk[:, 1, 1] * F32(height)
    cx = Kk[:, 0, 2] * F32(width)
    cy = Kk[:, 1, 2] * F32(height)
    fxo = Ks[:, 0, 0] * F32(width_source)
    fyo = Ks[:, 1, 1] * F32(height_source)
    cxo = Ks[:, 0, 2] * F32(width_source)
    cyo = Ks[:, 1, 2] * F32(height_source)
    r11, r12, r13, t1 = T[:, 0, 0], T[:, 0, 1], T[:, 0, 2], T[:, 0, 3]
    r21, r22, r23, t2 = T[:, 1, 0], T[:, 1, 1], T[:, 1, 2], T[:, 1, 3]
    r31, r32, r33, t3 = T[:, 2, 0], T[:, 2, 1], T[:, 2, 2], T[:, 2, 3]

    A = fxo * r11 + cxo * r31
    B = fxo * r12 + cxo * r32
    a = A / fx  # :262
    b = B / fy  # :263
    c = (-((cx * A) / fx) - ((cy * B) / fy)) + (fxo * r13 + cxo * r33)  # :264
    e = fxo * t1 + cxo * t3  # :265
    Fm = fyo * r21 + cyo * r31
    G = fyo * r22 + cyo * r32
    f = Fm / fx  # :267
    g = G / fy  # :268
    h = (-((cx * Fm) / fx) - ((cy * G) / fy)) + (fyo * r23 + cyo * r33)  # :269
    i = fyo * t2 + cyo * t3  # :270
    j = r31 / fx  # :272
    k = r32 / fy  # :273
    l = ((((-cx) * r31) / fx) - ((cy * r32) / fy)) + r33  # :274
    m = t3.copy()  # :275
    return dict(a=a, b=b, c=c, e=e, f=f, g=g, h=h, i=i, j=j, k=k, l=l, m=m)


COEFF_ORDER = ("a", "b", "c", "e", "f", "g", "h", "i", "j", "k", "l", "m")


def epipolar_coeffs(intrinsics_key, intrinsics_source, source_to_key_transform, height, width,
                    height_source, width_source):
    """planesweep_corr.py:228-300 (``EpipolarCoeffs.from_calib``).

    Returns ``(u_infs_h, v_infs_h, k_infs_h)`` as ``[N,1,H,W]`` and ``(m_u_h, m_v_h, m_k_h)`` as ``[N,1,1,1]``.
    """
    s = epipolar_scalars(intrinsics_key, intrinsics_source, source_to_key_transform, height, width,
                         height_source, width_source)
    x = (np.arange(width, dtype=np.float32) + F32(0.5))[None, None, :]  # :233-236
    y = (np.arange(height, dtype=np.float32) + F32(0.5))[None, :, None]  # :237

    def col(v):
        return v[:, None, None]

    u_infs_h = (col(s["a"]) * x + col(s["b"]) * y) + col(s["c"])  # :277-284
    v_infs_h = (col(s["f"]) * x + col(s["g"]) * y) + col(s["h"])  # :281-287
    k_infs_h = (col(s["j"]) * x + col(s["k"]) * y) + col(s["l"])  # :279-290
    r4 = lambda v: v[:, None, None, None].astype(np.float32)  # noqa: E731
    return (u_infs_h[:, None].astype(np.float32), v_infs_h[:, None].astype(np.float32),
            k_infs_h[:, None].astype(np.float32), r4(s["e"]), r4(s["i"]), r4(s["m"]))


# --------------------------------------------------------------------------------------
# a4 + a5: sampling points and geometric visibility mask
# --------------------------------------------------------------------------------------
def _replace_nonfinite(x):
    """planesweep_corr.py:336-338: +-inf -> +-1e9, nan -> 1e9."""
    x = x.copy()
    inf = np.isinf(x)
    x[inf] = F32(1e9) * np.sign(x[inf])
    x[np.isnan(x)] = F32(1e9)
    return x


def sampling_points(coeffs, invdepths):
    """planesweep_corr.py:333-349 and :489-512.

    ``coeffs``: output of ``epipolar_coeffs``; ``invdepths``: ``[N or 1, S, 1 or H, 1 or W]``.
    Returns ``us, vs, zs`` (``[N,S,H,W]`` fp32) and the geometric ``mask`` (bool).
    """
    u_h, v_h, k_h, m_u, m_v, m_k = coeffs
    ds = _f32(invdepths)
    with np.errstate(divide="ignore", invalid="ignore", over="ignore"):
        den = k_h + m_k * ds
        us = _replace_nonfinite((u_h + m_u * ds) / den)  # :334-338
        vs = _replace_nonfinite((v_h + m_v * ds) / den)  # :343-347
        zs = (_ONE / ds) * np.ones_like(us)  # :492, :497
        z_poles = -(m_k / k_h)  # :330
        visible_in_key = zs > 0  # :499
        visible_in_source = (((k_h > 0) & (zs > z_poles)) | ((k_h < 0) & (zs < z_poles))
                             | ((k_h == 0) & (m_k > 0)))  # :500-504
    mask = visible_in_key & visible_in_source  # :506
    return us.astype(np.float32), vs.astype(np.float32), zs.astype(np.float32), mask


# --------------------------------------------------------------------------------------
# a6: warp / grid_sample
# --------------------------------------------------------------------------------------
def grid_unnormalize(coord, size):
    """ATen ``grid_sampler_unnormalize`` for ``align_corners=False``.

    torch ``include/ATen/native/GridSampler.h:27-36`` states ``((coord + 1) * size - 1) / 2``; the
    shipped kernels contract the multiply-subtract, i.e. they evaluate
    ``fma(coord + 1, size / 2, -0.5)`` (probed against torch 2.11 CPU: 0 mismatches in 2e5 samples
    for the fused form, 0.3 % mismatching for the separately rounded one).
    """
    t = _f32(coord) + _ONE
    return _fma32(t, F32(size) / _TWO, F32(-0.5))


def bilinear_setup(us, vs, hs, ws, padding_mode="zeros"):
    """Sampling indices and weights of ``warp`` (planesweep_corr.py:85-93) + ATen bilinear grid_sample.

    Returns dict with integer corner ``ix0, iy0`` (int32, saturated to ``[-2, size]``), the four weights
    ``nw, ne, sw, se``, the four in-bounds flags and the zeros-padding sampling mask
    (planesweep_corr.py:99-102: sum of in-bounds weights >= 0.9999 -> 1 else 0).
    """
    with np.errstate(over="ignore", invalid="ignore"):
        xg = (_TWO * _f32(us)) / F32(ws) - _ONE  # :87
        yg = (_TWO * _f32(vs)) / F32(hs) - _ONE  # :88
        ix = grid_unnormalize(xg, ws)
        iy = grid_unnormalize(yg, hs)
        if padding_mode == "border":
            # ATen clip_coordinates (GridSampler.h): std::min(size - 1, std::max(coord, 0)); with the CPU's std::min / std::max a
            # nan coordinate comes out as size - 1
            ix = np.where(np.isnan(ix), F32(ws - 1), np.minimum(F32(ws - 1), np.maximum(ix, F32(0)))).astype(np.float32)
            iy = np.where(np.isnan(iy), F32(hs - 1), np.minimum(F32(hs - 1), np.maximum(iy, F32(0)))).astype(np.float32)
        x0 = np.floor(ix)
        y0 = np.floor(iy)
        wx1 = ix - x0  # "w"
        wx0 = _ONE - wx1  # "e"
        wy1 = iy - y0  # "n"
        wy0 = _ONE - wy1  # "s"
        nw = wy0 * wx0
        ne = wy0 * wx1
        sw = wy1 * wx0
        se = wy1 * wx1
    x0_in = (x0 >= 0) & (x0 <= ws - 1)
    x1_in = (x0 + 1 >= 0) & (x0 + 1 <= ws - 1)
    y0_in = (y0 >= 0) & (y0 <= hs - 1)
    y1_in = (y0 + 1 >= 0) & (y0 + 1 <= hs - 1)
    in_nw, in_ne, in_sw, in_se = y0_in & x0_in, y0_in & x1_in, y1_in & x0_in, y1_in & x1_in
    z = F32(0)
    msum = ((np.where(in_nw, nw, z) + np.where(in_ne, ne, z)) + np.where(in_sw, sw, z)) + np.where(in_se, se, z)
    samp_mask = (~(msum < F32(0.9999))) & (msum > 0)  # :101-102 (nan cannot occur: us/vs are finite)
    if padding_mode == "border":
        samp_mask = np.ones_like(samp_mask)           # :95-96: border padding -> the mask is all ones
    ix0 = np.clip(x0, -2, ws).astype(np.int32)
    iy0 = np.clip(y0, -2, hs).astype(np.int32)
    return dict(ix0=ix0, iy0=iy0, nw=nw, ne=ne, sw=sw, se=se, in_nw=in_nw, in_ne=in_ne, in_sw=in_sw,
                in_se=in_se, samp_mask=samp_mask.astype(np.float32))


def _l2_normalize(x, axis):
    """planesweep_corr.py:8-10."""
    norm = np.sqrt(np.sum(x.astype(np.float32) ** 2, axis=axis, keepdims=True, dtype=np.float32))
    return (x / (norm + F32(1e-9))).astype(np.float32)


# --------------------------------------------------------------------------------------
# a7: TorchCorr
# --------------------------------------------------------------------------------------
def torch_corr(feat_ref, feat_src, us, vs, mask=None, normalize="dim", return_debug=False, padding_mode="zeros"):
    """planesweep_corr.py:152-195 (``TorchCorr.forward``, zeros or border padding), the reference's own algorithm:
    all-pairs correlation matrix, then bilinear sampling of that matrix at the S epipolar points.

    feat_ref ``[N,C,h,w]``, feat_src ``[N,C,hs,ws]``, us/vs ``[N,S,h,w]`` (= grids[:, :, 0/1]),
    mask ``[N,S,h,w]`` bool or None.  Returns ``corr, corr_mask`` both ``[N,S,h,w]`` fp32.
    """
    feat_ref = _f32(feat_ref)
    feat_src = _f32(feat_src)
    N, C, h, w = feat_ref.shape
    hs, ws = feat_src.shape[-2:]
    S = us.shape[1]
    fr = feat_ref.reshape(N, C, h * w)
    fs = feat_src.reshape(N, C, hs * ws)
    if normalize and normalize != "dim":  # :170-172
        fr = _l2_normalize(fr, 1)
        fs = _l2_normalize(fs, 1)
    b = bilinear_setup(us, vs, hs, ws, padding_mode)
    corr = np.empty((N, S, h, w), np.float32)
    P = h * w
    for n in range(N):
        allp = np.matmul(fr[n].T, fs[n])  # :174-176  [h*w, hs*ws]
        if normalize == "dim":
            allp = allp / np.sqrt(F32(C))  # :178-179
        flat = allp.reshape(-1)
        base = (np.arange(P, dtype=np.int64) * (hs * ws)).reshape(1, h, w)
        x0 = np.clip(b["ix0"][n], 0, ws - 1).astype(np.int64)
        x1 = np.clip(b["ix0"][n] + 1, 0, ws - 1).astype(np.int64)
        y0 = np.clip(b["iy0"][n], 0, hs - 1).astype(np.int64)
        y1 = np.clip(b["iy0"][n] + 1, 0, hs - 1).astype(np.int64)
        z = F32(0)
        acc = np.where(b["in_nw"][n], flat[base + y0 * ws + x0], z) * b["nw"][n]
        acc = acc + np.where(b["in_ne"][n], flat[base + y0 * ws + x1], z) * b["ne"][n]
        acc = acc + np.where(b["in_sw"][n], flat[base + y1 * ws + x0], z) * b["sw"][n]
        acc = acc + np.where(b["in_se"][n], flat[base + y1 * ws + x1], z) * b["se"][n]
        corr[n] = acc
    corr_mask = b["samp_mask"]
    corr = corr * corr_mask  # :189
    if mask is not None:  # :191-193
        mf = mask.astype(np.float32)
        corr = corr * mf
        corr_mask = corr_mask * mf
    if return_debug:
        return corr, corr_mask, b
    return corr, corr_mask


# --------------------------------------------------------------------------------------
# a8: WarpOnlyCorr
# --------------------------------------------------------------------------------------
def grid_sample_bilinear_zeros(x, b):
    """Bilinear gather of ``x [N,C,hs,ws]`` at the points described by ``b`` (``bilinear_setup`` output with
    leading dims ``[N, ...]``).  Returns ``[N, C, ...]``."""
    x = _f32(x)
    N, C, hs, ws = x.shape
    out_shape = b["ix0"].shape[1:]
    out = np.empty((N, C) + out_shape, np.float32)
    z = F32(0)
    for n in range(N):
        x0 = np.clip(b["ix0"][n], 0, ws - 1).astype(np.int64)
        x1 = np.clip(b["ix0"][n] + 1, 0, ws - 1).astype(np.int64)
        y0 = np.clip(b["iy0"][n], 0, hs - 1).astype(np.int64)
        y1 = np.clip(b["iy0"][n] + 1, 0, hs - 1).astype(np.int64)
        img = x[n].reshape(C, hs * ws)
        acc = np.where(b["in_nw"][n], img[:, y0 * ws + x0], z) * b["nw"][n]
        acc = acc + np.where(b["in_ne"][n], img[:, y0 * ws + x1], z) * b["ne"][n]
        acc = acc + np.where(b["in_sw"][n], img[:, y1 * ws + x0], z) * b["sw"][n]
        acc = acc + np.where(b["in_se"][n], img[:, y1 * ws + x1], z) * b["se"][n]
        out[n] = acc
    return out


def warp_only_corr(feat_src, us, vs, normalize=False):
    """planesweep_corr.py:107-140 + ``warp_multi`` :13-46 (zeros padding).

    Returns warped ``[N,S,C,h,w]`` and warping mask ``[N,S,h,w]``; the geometric mask is ignored (:116-140).
    """
    feat_src = _f32(feat_src)
    hs, ws = feat_src.shape[-2:]
    if normalize == "before":  # :128-129
        feat_src = _l2_normalize(feat_src, 1)
    b = bilinear_setup(us, vs, hs, ws)
    warped = grid_sample_bilinear_zeros(feat_src, b)  # [N,C,S,h,w]
    warped = np.ascontiguousarray(np.moveaxis(warped, 1, 2))  # [N,S,C,h,w]
    if normalize and normalize != "before" and normalize != "dim":  # :135-136
        warped = _l2_normalize(warped, 2)
    m = b["samp_mask"]
    warped = warped * m[:, :, None]  # :138
    return warped.astype(np.float32), m


# --------------------------------------------------------------------------------------
# a9: the block
# --------------------------------------------------------------------------------------
def planesweep_correlation(feat_key, intrinsics_key, feat_sources, source_to_key_transforms,
                           intrinsics_sources=None, num_sampling_points=None, min_depth=None, max_depth=None,
                           sampling_invdepths=None, sampling_type="linear_invdepth", warp_only=False,
                           normalize="dim", return_debug=False):
    """planesweep_corr.py:396-427 (``PlanesweepCorrelation.forward``).

    Returns ``(corrs, masks, sampling_invdepths)`` like the reference; with ``return_debug`` a fourth item, a
    list of per-view dicts holding ``us, vs, zs, geom_mask, ix0, iy0``.
    """
    feat_key = _f32(feat_key)
    N, C, h, w = feat_key.shape
    if intrinsics_sources is None:  # :441-442
        intrinsics_sources = [intrinsics_key] * len(feat_sources)
    assert len(feat_sources) == len(source_to_key_transforms) == len(intrinsics_sources)  # :444
    if min_depth is not None and max_depth is not None:  # :473-480
        assert sampling_invdepths is None and num_sampling_points is not None and sampling_type is not None
        inv = compute_sampling_invdepths(min_depth, max_depth, num_sampling_points, sampling_type)
    else:
        assert num_sampling_points is None and min_depth is None and max_depth is None
        assert sampling_invdepths is not None
        inv = _f32(sampling_invdepths)
    assert inv.ndim >= 2  # :483
    while inv.ndim < 4:  # :484-485
        inv = inv[..., None]
    corrs, masks, dbg = [], [], []
    for v, T in enumerate(source_to_key_transforms):
        fs = _f32(feat_sources[v])
        hs, ws = fs.shape[-2:]
        co = epipolar_coeffs(intrinsics_key, intrinsics_sources[v], T, h, w, hs, ws)
        us, vs, zs, gmask = sampling_points(co, inv)
        if warp_only:
            c, m = warp_only_corr(fs, us, vs, normalize=normalize)
            b = bilinear_setup(us, vs, hs, ws) if return_debug else None
        else:
            c, m, b = torch_corr(feat_key, fs, us, vs, mask=gmask, normalize=normalize, return_debug=True)
        corrs.append(c)
        masks.append(m)
        if return_debug:
            dbg.append(dict(us=us, vs=vs, zs=zs, geom_mask=gmask, ix0=b["ix0"], iy0=b["iy0"]))
    if return_debug:
        return corrs, masks, inv, dbg
    return corrs, masks, inv


# --------------------------------------------------------------------------------------
# f-4: backward of the correlation with respect to the feature maps
# --------------------------------------------------------------------------------------
def _l2_normalize_backward(x, g, axis):
    """Gradient of ``x / (||x|| + 1e-9)`` (planesweep_corr.py:8-10) for upstream gradient ``g``."""
    x = x.astype(np.float64)
    g = g.astype(np.float64)
    r = np.sqrt(np.sum(x * x, axis=axis, keepdims=True))
    den = r + 1e-9
    dot = np.sum(g * x, axis=axis, keepdims=True)
    with np.errstate(divide="ignore", invalid="ignore"):
        corr = np.where(r > 0, x * dot / (den * den * r), 0.0)
    return g / den - corr


def planesweep_correlation_backward(feat_key, intrinsics_key, feat_sources, source_to_key_transforms, grad_corrs,
                                    intrinsics_sources=None, num_sampling_points=None, min_depth=None, max_depth=None,
                                    sampling_invdepths=None, sampling_type="linear_invdepth", normalize="dim"):
    """What autograd returns for ``sum_v <grad_corrs[v], corrs[v]>`` through the reference block with
    ``warp_only=False`` (``TorchCorr.forward``, planesweep_corr.py:152-195: matmul :174, ``/ sqrt(C)`` :178-179,
    ``grid_sample`` :91, ``* corr_mask`` :189, ``* mask`` :191-192).  Sampling points carry no gradient (:436, :464,
    :489 are ``no_grad``).  Returns ``(grad_key [N,C,h,w], [grad_src_v [N,C,hs,ws]])`` as float32 (float64 sums).
    """
    feat_key = _f32(feat_key)
    N, C, h, w = feat_key.shape
    if intrinsics_sources is None:
        intrinsics_sources = [intrinsics_key] * len(feat_sources)
    if min_depth is not None and max_depth is not None:
        inv = compute_sampling_invdepths(min_depth, max_depth, num_sampling_points, sampling_type)
    else:
        inv = _f32(sampling_invdepths)
    while inv.ndim < 4:
        inv = inv[..., None]
    l2 = bool(normalize) and normalize != "dim"
    scale = np.float64(np.sqrt(F32(C))) if normalize == "dim" else np.float64(1.0)
    P = h * w
    key_in = feat_key.reshape(N, C, P)
    key = _l2_normalize(key_in, 1) if l2 else key_in
    gkey = np.zeros((N, C, P), np.float64)
    gsrcs = []
    for v, T in enumerate(source_to_key_transforms):
        fs_in = _f32(feat_sources[v])
        hs, ws = fs_in.shape[-2:]
        fs_in = fs_in.reshape(N, C, hs * ws)
        fs = _l2_normalize(fs_in, 1) if l2 else fs_in
        co = epipolar_coeffs(intrinsics_key, intrinsics_sources[v], T, h, w, hs, ws)
        us, vs, _, gmask = sampling_points(co, inv)
        b = bilinear_setup(us, vs, hs, ws)
        valid = b["samp_mask"].astype(np.float64) * gmask.astype(np.float64)      # [N,S,h,w]
        g = _f32(grad_corrs[v]).astype(np.float64) * valid / scale
        gs = np.zeros((N, hs * ws, C), np.float64)
        S = g.shape[1]
        for n in range(N):
            x0 = np.clip(b["ix0"][n], 0, ws - 1).astype(np.int64)
            x1 = np.clip(b["ix0"][n] + 1, 0, ws - 1).astype(np.int64)
            y0 = np.clip(b["iy0"][n], 0, hs - 1).astype(np.int64)
            y1 = np.clip(b["iy0"][n] + 1, 0, hs - 1).astype(np.int64)
            key_pc = key[n].T.astype(np.float64)                                  # [P,C]
            src_tc = fs[n].T.astype(np.float64)                                   # [hs*ws,C]
            for idx, wname, iname in ((y0 * ws + x0, "nw", "in_nw"), (y0 * ws + x1, "ne", "in_ne"),
                                      (y1 * ws + x0, "sw", "in_sw"), (y1 * ws + x1, "se", "in_se")):
                with np.errstate(invalid="ignore"):
                    a = np.where(b[iname][n] & (valid[n] > 0), b[wname][n].astype(np.float64) * g[n], 0.0)  # [S,h,w]
                for s in range(S):
                    a_s = a[s].reshape(P)
                    nz = np.nonzero(a_s)[0]
                    if nz.size == 0:
                        continue
                    t = idx[s].reshape(P)[nz]
                    gkey[n][:, nz] += (src_tc[t] * a_s[nz, None]).T
                    np.add.at(gs[n], t, key_pc[nz] * a_s[nz, None])
        gs = np.ascontiguousarray(np.moveaxis(gs, 2, 1))                          # [N,C,hs*ws]
        if l2:
            gs = _l2_normalize_backward(fs_in, gs, 1)
        gsrcs.append(gs.reshape(N, C, hs, ws).astype(np.float32))
    if l2:
        gkey = _l2_normalize_backward(key_in, gkey, 1)
    return gkey.reshape(N, C, h, w).astype(np.float32), gsrcs
