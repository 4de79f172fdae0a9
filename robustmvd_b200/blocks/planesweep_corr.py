"""Drop-in for ``rmvd/models/blocks/planesweep_corr.py`` on top of the sm_100a C ABI.

Same names, constructor and ``forward`` signatures, return values and assertions as the reference
(``PlanesweepCorrelation`` :371-521, ``TorchCorr`` :143-195, ``WarpOnlyCorr`` :107-140, ``warp`` :49-104,
``warp_multi`` :13-46, ``compute_sampling_invdepths`` :524-555).  All arithmetic runs in the CUDA library
(``include/rmvd_b200.h``); there is no PyTorch or CPU fallback.  When gradients are enabled and a feature map requires
them, the correlation outputs carry an autograd node whose backward is the CUDA kernel behind ``rmvd_psweep_corr_bwd``
(SURVEY.md row f-4); sampling points carry no gradient, like in the reference (:436, :464, :489 are ``no_grad``).
"""
import numpy as np
import ctypes

import torch
import torch.nn as nn

from .. import _lib

NORM_NONE, NORM_DIM, NORM_BEFORE, NORM_AFTER = 0, 1, 2, 3


def _to_torch(x, device=None):
    if isinstance(x, np.ndarray):
        x = torch.from_numpy(x)
    if isinstance(x, torch.Tensor) and device is not None:
        x = x.to(device)
    return x


def _f32c(t):
    return t.detach().to(torch.float32).contiguous()


def _calib(t, N, rows, device, what):
    """Calibration tensor for the C ABI: float32, contiguous, exactly ``[N, rows, rows]``.  The kernels index these arrays as
    ``K + n * 9`` / ``T + n * 16`` and only see pointers, so shapes are settled here: the reference slices ``[:, i, j]`` and thus also
    takes ``[N,3,4]`` transforms (padded with the row 0 0 0 1) and batch-1 tensors that broadcast over N (expanded)."""
    t = _f32c(_to_torch(t, device))
    if t.ndim == 2:
        t = t[None]
    if t.ndim != 3:
        raise _lib.RmvdError(f"{what}: expected [N,{rows},{rows}], got {tuple(t.shape)}")
    if rows == 4 and tuple(t.shape[1:]) == (3, 4):
        last = torch.tensor([0.0, 0.0, 0.0, 1.0], dtype=torch.float32, device=t.device).expand(t.shape[0], 1, 4)
        t = torch.cat([t, last], 1)
    if tuple(t.shape[1:]) != (rows, rows):
        raise _lib.RmvdError(f"{what}: expected [N,{rows},{rows}], got {tuple(t.shape)}")
    if t.shape[0] == 1 and N > 1:
        t = t.expand(N, rows, rows)
    if t.shape[0] != N:
        raise _lib.RmvdError(f"{what}: batch {t.shape[0]} does not match the feature batch {N}")
    return t.contiguous()


def _corr_norm_mode(normalize):
    """TorchCorr semantics (:170-179): "dim" -> /sqrt(C); True / "before" -> L2-normalise both inputs."""
    if normalize == "dim":
        return NORM_DIM
    return NORM_BEFORE if normalize else NORM_NONE


def _warp_norm_mode(normalize):
    """WarpOnlyCorr semantics (:128-136): "before" -> normalise source; True / "after" -> normalise warped."""
    if normalize == "before":
        return NORM_BEFORE
    if normalize and normalize != "dim":
        return NORM_AFTER
    return NORM_NONE


def _workspace(nbytes, device):
    return torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=device)


def compute_sampling_invdepths(min_depth, max_depth, num_samples, sampling_type="linear_invdepth"):
    """planesweep_corr.py:524-555: ``[1 or N, S]`` inverse depths (host-side schedule, S values per sample).

    Evaluated on the HOST with the reference's own operation order (S values per sample): torch's CUDA division
    by a scalar multiplies by the reciprocal, which would move the last bit of the schedule.
    """
    if isinstance(min_depth, (np.float32, float, int)):
        min_depth = torch.tensor([float(min_depth)])
    if isinstance(max_depth, (np.float32, float, int)):
        max_depth = torch.tensor([float(max_depth)])
    min_depth = _to_torch(min_depth).cpu()[..., None]
    max_depth = _to_torch(max_depth).cpu()[..., None]
    min_invdepth = 1 / max_depth
    max_invdepth = 1 / min_depth
    steps = torch.arange(0, num_samples, dtype=min_invdepth.dtype, device=min_invdepth.device)[None, ...]
    if sampling_type == "linear_invdepth":
        return min_invdepth + steps * (max_invdepth - min_invdepth) / (num_samples - 1)
    elif sampling_type == "linear_depth":
        inv = 1 / (min_depth + steps * (max_depth - min_depth) / (num_samples - 1))
        return inv.flip(1)
    raise ValueError(f"unknown sampling_type {sampling_type}")


def _inv_strides(inv, N):
    """Element strides (n, s, y, x) of a contiguous 4-D inverse-depth tensor with broadcast dims -> 0."""
    assert inv.ndim == 4
    st = list(inv.stride())
    for k in range(4):
        if inv.shape[k] == 1:
            st[k] = 0
    assert inv.shape[0] in (1, N), "sampling_invdepths batch must be 1 or N"
    return st


def _pixel_centres(h_ref, w_ref, device):
    yy, xx = torch.meshgrid(torch.arange(h_ref, device=device), torch.arange(w_ref, device=device), indexing="ij")
    return torch.stack((xx + 0.5, yy + 0.5), 0).float()          # 2, h_ref, w_ref  (:68-77)


def _warp_launch(x, grid, K, x_repeat, padding_mode):
    """``x [K / x_repeat, C, hs, ws]``, ``grid [K, 2, ho, wo]`` -> ``out [K, C, ho, wo]``, ``mask [K, 1, ho, wo]``."""
    C, hs, ws = x.shape[1:]
    ho, wo = grid.shape[-2:]
    out = torch.empty((K, C, ho, wo), dtype=torch.float32, device=x.device)
    mask = torch.empty((K, 1, ho, wo), dtype=torch.float32, device=x.device)
    lib = _lib.load()
    step = max(x_repeat, 65535 // x_repeat * x_repeat)        # gridDim.y carries K: chunk very large K on image boundaries
    for k0 in range(0, K, step):
        k1 = min(K, k0 + step)
        st = lib.rmvd_warp_grid_fwd(_lib.ptr(x[k0 // x_repeat:]), _lib.ptr(grid[k0:k1]), k1 - k0, C, hs, ws, ho, wo,
                                    1 if padding_mode == "border" else 0, x_repeat, _lib.ptr(out[k0:k1]),
                                    _lib.ptr(mask[k0:k1]), _lib.stream_ptr())
        _lib.check(st, "rmvd_warp_grid_fwd")
    return out, mask


def warp(x, offset=None, grid=None, padding_mode="border"):
    """planesweep_corr.py:49-104.  ``x [N,C,H,W]``; ``grid [N,2,h_out,w_out]`` in source pixels (or ``offset``
    relative to pixel centres).  Returns ``(output [N,C,h_out,w_out], mask [N,1,h_out,w_out])``."""
    assert (offset is None and grid is not None) or (grid is None and offset is not None)
    assert padding_mode in ("border", "zeros")
    if offset is not None:
        grid = _pixel_centres(offset.shape[-2], offset.shape[-1], x.device).unsqueeze(0) + offset
    x = _f32c(x)
    return _warp_launch(x, _f32c(grid), x.shape[0], 1, padding_mode)


def warp_multi(x, offsets=None, grids=None, padding_mode="border"):
    """planesweep_corr.py:13-46: ``x [N,C,hs,ws]``, ``grids [N,S,2,h,w]`` -> ``[N,S,C,h,w]``, ``[N,S,h,w]`` in ONE launch:
    the S-fold repeat of the source (:33) is not materialised, grid ``n * S + s`` samples image ``n``."""
    assert (offsets is None and grids is not None) or (grids is None and offsets is not None)
    assert padding_mode in ("border", "zeros")
    loc = offsets if offsets is not None else grids
    N, S, _, H, W = loc.shape
    if offsets is not None:
        loc = _pixel_centres(H, W, x.device).view(1, 1, 2, H, W) + offsets
    x = _f32c(x)
    out, mask = _warp_launch(x, _f32c(loc).view(N * S, 2, H, W), N * S, S, padding_mode)
    return out.view(N, S, x.shape[1], H, W), mask.view(N, S, H, W)


class WarpOnlyCorr(nn.Module):
    """planesweep_corr.py:107-140 on caller-supplied grids (zeros or border padding)."""

    def __init__(self, normalize=False, padding_mode="zeros"):
        super().__init__()
        self.normalize = normalize
        self.padding_mode = padding_mode
        if isinstance(self.normalize, str):
            assert self.normalize in ["dim", "before", "after"]

    def forward(self, feat_ref, feat_src, grids=None, mask=None):
        if self.normalize == "before":
            feat_src = feat_src / (torch.linalg.norm(feat_src, dim=1, keepdim=True) + 1e-9)
        warped, warping_mask = warp_multi(x=feat_src, grids=grids, padding_mode=self.padding_mode)
        if self.normalize and self.normalize != "before" and self.normalize != "dim":
            warped = warped / (torch.linalg.norm(warped, dim=2, keepdim=True) + 1e-9)
        return warped * warping_mask.unsqueeze(2), warping_mask


class TorchCorr(nn.Module):
    """planesweep_corr.py:143-195 on caller-supplied grids: ``rmvd_corr_grid_fwd`` (zeros or border padding)."""

    def __init__(self, normalize=False, padding_mode="zeros"):
        super().__init__()
        self.normalize = normalize
        self.padding_mode = padding_mode
        if isinstance(self.normalize, str):
            assert self.normalize in ["dim", "before"]
        assert padding_mode in ("zeros", "border")

    def forward(self, feat_ref, feat_src, grids=None, mask=None):
        feat_ref, feat_src, grids = _f32c(feat_ref), _f32c(feat_src), _f32c(grids)
        N, C, h, w = feat_ref.shape
        S = grids.shape[1]
        hs, ws = feat_src.shape[-2:]
        dev = feat_ref.device
        m8 = None if mask is None else (mask != 0).to(torch.uint8).expand(N, S, h, w).contiguous()
        corr = torch.empty((N, S, h, w), dtype=torch.float32, device=dev)
        cmask = torch.empty_like(corr)
        lib = _lib.load()
        nb = lib.rmvd_psweep_workspace_bytes(N, C, h, w, 1, _lib.int_array([hs]), _lib.int_array([ws]))
        wsp = _workspace(nb, dev)
        st = lib.rmvd_corr_grid_fwd(_lib.ptr(feat_ref), _lib.ptr(feat_src), _lib.ptr(grids), _lib.ptr(m8), N, C, h, w,
                                    hs, ws, S, _corr_norm_mode(self.normalize), 1 if self.padding_mode == "border" else 0,
                                    _lib.ptr(corr), _lib.ptr(cmask), _lib.ptr(wsp), wsp.numel(), _lib.stream_ptr())
        _lib.check(st, "rmvd_corr_grid_fwd")
        return corr, cmask


_INVDEPTH_CACHE = {}


class _WarpFunction(torch.autograd.Function):
    """Autograd node of one view of the warp-only block (``warp_only=True``, normalisation off): forward = the block's own
    ``rmvd_psweep_warp_fwd`` call, backward = ``rmvd_psweep_warp_bwd``; the 0/1 mask output is not differentiable."""

    @staticmethod
    def forward(ctx, block, args, view, feat_src):
        corrs, masks, inv = block._forward_impl(feat_key=args["feat_key"], feat_sources=[feat_src], norm_mode=NORM_NONE,
                                                intrinsics_key=args["intrinsics_key"],
                                                source_to_key_transforms=[args["source_to_key_transforms"][view]],
                                                intrinsics_sources=None if args["intrinsics_sources"] is None else [args["intrinsics_sources"][view]],
                                                **args["sampling"])
        ctx.block, ctx.args, ctx.view, ctx.inv = block, args, view, inv
        block._inv_last = inv
        ctx.src_shape = tuple(feat_src.shape)
        ctx.mark_non_differentiable(masks[0])
        return corrs[0], masks[0]

    @staticmethod
    def backward(ctx, grad_warped, _grad_mask):
        a, v = ctx.args, ctx.view
        Ks = a["intrinsics_key"] if a["intrinsics_sources"] is None else a["intrinsics_sources"][v]
        return (None, None, None, ctx.block._warp_backward_impl(grad_warped, ctx.src_shape, a["feat_key"].shape, a["intrinsics_key"], Ks,
                                                                a["source_to_key_transforms"][v], ctx.inv, a["sampling"].get("plane_range")))


class _CorrFunction(torch.autograd.Function):
    """Autograd node of the correlation (``warp_only=False``): forward = ``rmvd_psweep_corr_fwd``, backward =
    ``rmvd_psweep_corr_bwd``; the 0/1 mask output is not differentiable."""

    @staticmethod
    def forward(ctx, block, args, mode, feat_key, *feat_sources):
        _, _, inv = block._forward_impl(feat_key=feat_key, feat_sources=list(feat_sources), norm_mode=mode, **args)
        corr, mask = block._stacked
        block._stacked = None
        block._inv = inv
        ctx.block, ctx.args, ctx.mode, ctx.inv = block, args, mode, inv
        ctx.save_for_backward(feat_key, *feat_sources)
        ctx.mark_non_differentiable(mask)
        return corr, mask

    @staticmethod
    def backward(ctx, grad_corr, _grad_mask):
        feat_key, *feat_sources = ctx.saved_tensors
        a = ctx.args
        need = ctx.needs_input_grad[3:]
        gkey, gsrcs = ctx.block._backward_impl(
            grad_corr, feat_key, feat_sources, a["intrinsics_key"], a["source_to_key_transforms"],
            a["intrinsics_sources"], ctx.inv, a["plane_range"], ctx.mode, need[0], list(need[1:]))
        return (None, None, None, gkey, *gsrcs)


class _LazyView:
    """Per-source-view state of the reference block (planesweep_corr.py:389-394), materialised on first access by
    ``rmvd_sampling_points_fwd``: the fused correlation never writes sampling points or coefficients."""

    def __init__(self, calib):
        self._calib = calib          # (Kk, Ks_v, T_v, inv_c, N, h, w, hs, ws, S)
        self._pts = None
        self._co = None

    def _run(self, points):
        Kk, Ks, T, inv_c, N, h, w, hs, ws, S = self._calib
        dev = Kk.device
        ist = _lib.i64_array(_inv_strides(inv_c, N))
        us = vs = zs = mask = pix = mvec = None
        if points:
            us = torch.empty((N, S, h, w), dtype=torch.float32, device=dev)
            vs, zs = torch.empty_like(us), torch.empty_like(us)
            mask = torch.empty((N, S, h, w), dtype=torch.uint8, device=dev)
        else:
            pix = torch.empty((N, 3, h, w), dtype=torch.float32, device=dev)
            mvec = torch.empty((N, 3), dtype=torch.float32, device=dev)
        st = _lib.load().rmvd_sampling_points_fwd(_lib.ptr(Kk), _lib.ptr(Ks), _lib.ptr(T), _lib.ptr(inv_c), ist, N, h, w, hs, ws, S,
                                                  _lib.ptr(us), _lib.ptr(vs), _lib.ptr(zs), _lib.ptr(mask), _lib.ptr(pix), _lib.ptr(mvec),
                                                  _lib.stream_ptr())
        _lib.check(st, "rmvd_sampling_points_fwd")
        if points:
            self._pts = (us, vs, zs, mask.bool())
        else:
            self._co = (pix, mvec)

    def points(self):
        if self._pts is None:
            self._run(True)
        return self._pts

    def coeffs(self):
        if self._co is None:
            self._run(False)
        return self._co


class EpipolarSamplingPoints:
    """planesweep_corr.py:352-368: ``us``, ``vs``, ``zs`` float32 and ``mask`` bool, all ``[N,S,h,w]`` (lazily computed)."""

    def __init__(self, lazy):
        self._lazy = lazy
        _, _, _, _, _, h, w, _, _, S = lazy._calib
        self.height, self.width, self._num = h, w, S

    us = property(lambda self: self._lazy.points()[0])
    vs = property(lambda self: self._lazy.points()[1])
    zs = property(lambda self: self._lazy.points()[2])
    mask = property(lambda self: self._lazy.points()[3])

    @property
    def num(self):
        return self._num

    def get_uvs(self):
        return torch.stack((self.us, self.vs), 2)  # (N, S, 2, H, W)


class EpipolarCoeffs:
    """planesweep_corr.py:198-331 (the attributes and derived properties; ``from_calib`` runs inside the kernels)."""

    def __init__(self, lazy):
        self._lazy = lazy
        _, _, _, _, _, h, w, _, _, _ = lazy._calib
        self.height, self.width, self.eps = h, w, 1e-9
        self._u_infs = self._v_infs = self._z_poles = self._epipole = None

    u_infs_h = property(lambda self: self._lazy.coeffs()[0][:, 0:1])
    v_infs_h = property(lambda self: self._lazy.coeffs()[0][:, 1:2])
    k_infs_h = property(lambda self: self._lazy.coeffs()[0][:, 2:3])
    m_u_h = property(lambda self: self._lazy.coeffs()[1][:, 0].reshape(-1, 1, 1, 1))
    m_v_h = property(lambda self: self._lazy.coeffs()[1][:, 1].reshape(-1, 1, 1, 1))
    m_k_h = property(lambda self: self._lazy.coeffs()[1][:, 2].reshape(-1, 1, 1, 1))

    @staticmethod
    def _finite(x):                                   # :303-304, :311-312
        x = x.clone()
        bad = torch.isinf(x)
        x[bad] = 1e9 * torch.sign(x)[bad]
        return x

    @property
    def u_infs(self):
        if self._u_infs is None:
            self._u_infs = self._finite(self.u_infs_h / self.k_infs_h)
        return self._u_infs

    @property
    def v_infs(self):
        if self._v_infs is None:
            self._v_infs = self._finite(self.v_infs_h / self.k_infs_h)
        return self._v_infs

    @property
    def epipole(self):
        if self._epipole is None:
            self._epipole = (self.m_u_h / self.m_k_h, self.m_v_h / self.m_k_h)
        return self._epipole

    @property
    def z_poles(self):
        if self._z_poles is None:
            self._z_poles = -(self.m_k_h / self.k_infs_h)
        return self._z_poles


class PlanesweepCorrelation(nn.Module):
    """planesweep_corr.py:371-521.  ``forward`` returns ``(corrs, masks, sampling_invdepths)`` like the reference.

    ``plane_range=(s_begin, s_end)`` (extension, default all planes) restricts the call to a depth slab; outputs
    then hold only those planes.  ``debug=True`` additionally keeps ``self.debug`` with the integer
    sampling indices and the geometric mask of every view.  ``out=(corr, mask)`` (extension): caller-provided contiguous
    float32 device tensors ``[V, N, planes, h, w]`` that receive the correlation outputs instead of fresh allocations
    (``planes`` = the slab's length, or S: the slab is then written at its own plane offset of the full-height tensors).
    """

    @torch.no_grad()
    def __init__(self, warp_only=False, normalize="dim"):
        super().__init__()
        self.warp_only = warp_only
        self.normalize = normalize
        if isinstance(normalize, str):
            assert normalize in (["dim", "before", "after"] if warp_only else ["dim", "before"])
        self.corrs = []
        self.masks = []
        self.debug = None
        self.coeffs = []               # per source view, like the reference's attributes of the same names (:389-391),
        self.sampling_points = []      # computed on first read of an attribute (rmvd_sampling_points_fwd)

    def reset(self):
        self.corrs = []
        self.masks = []
        self.debug = None
        self.coeffs = []
        self.sampling_points = []

    @torch.no_grad()
    def get_sampling_invdepths(self, num_sampling_points=None, min_depth=None, max_depth=None,
                               sampling_invdepths=None, sampling_type="linear_invdepth", device=None):
        if min_depth is not None and max_depth is not None:  # :473-480
            assert sampling_invdepths is None
            assert num_sampling_points is not None
            assert sampling_type is not None
            # the schedule of a (range, count, type) is a constant: keep its device copy (a fresh pageable H2D copy per call
            # would block the host on the launching stream and serialise a pipelined caller)
            if not (isinstance(min_depth, (int, float)) and isinstance(max_depth, (int, float))):   # per-sample ranges
                return self._as_4d(_to_torch(compute_sampling_invdepths(min_depth, max_depth, num_sampling_points,
                                                                        sampling_type)).to(device))
            key = (float(min_depth), float(max_depth), int(num_sampling_points), sampling_type, str(device))
            cached = _INVDEPTH_CACHE.get(key)
            if cached is None:
                cached = self._as_4d(_to_torch(compute_sampling_invdepths(min_depth, max_depth, num_sampling_points,
                                                                         sampling_type)).to(device))
                if cached.is_cuda:                       # the pageable H2D copy ran on the stream current NOW; later calls clone the
                    torch.cuda.current_stream(cached.device).synchronize()   # entry from other streams: make it globally complete once
                if len(_INVDEPTH_CACHE) >= 16:
                    _INVDEPTH_CACHE.clear()
                _INVDEPTH_CACHE[key] = cached
            return cached.clone()   # device-side copy: the caller owns the returned tensor, like in the reference
        else:
            assert num_sampling_points is None and min_depth is None and max_depth is None
            assert sampling_invdepths is not None
        return self._as_4d(_to_torch(sampling_invdepths).to(device))

    @staticmethod
    def _as_4d(sampling_invdepths):
        assert sampling_invdepths.ndim >= 2  # (N, S) or (N, S, H) or (N, S, H, W)
        while sampling_invdepths.ndim < 4:
            sampling_invdepths = sampling_invdepths.unsqueeze(-1)
        return sampling_invdepths

    def forward(self, feat_key, intrinsics_key, feat_sources, source_to_key_transforms, intrinsics_sources=None,
                num_sampling_points=None, min_depth=None, max_depth=None, sampling_invdepths=None,
                sampling_type="linear_invdepth", plane_range=None, debug=False, out=None, fused_leaky_relu=None,
                key_activated_out=None):
        """The reference's signature (planesweep_corr.py:396-427) plus this library's extensions ``plane_range``, ``debug``, ``out``
        and ``fused_leaky_relu=slope``: ``feat_key`` / ``feat_sources`` are then the PRE-activation outputs of the encoder's last
        convolution and LeakyReLU(slope) is applied inside the kernel's operand emission (rmvd_psweep_corr_fwd_act);
        ``key_activated_out`` ([N,C,h,w], not overlapping ``feat_key``) receives the activated key map.  Forward only."""
        needs_grad = torch.is_grad_enabled() and not self.warp_only and len(feat_sources) > 0 and (
            feat_key.requires_grad or any(f.requires_grad for f in feat_sources))
        if fused_leaky_relu is not None:
            if needs_grad or self.warp_only or debug:
                raise _lib.RmvdError("fused_leaky_relu: forward-only correlation path (no grad, no warp_only, no debug outputs)")
            return self._forward_impl(feat_key, intrinsics_key, feat_sources, source_to_key_transforms,
                                      intrinsics_sources, num_sampling_points, min_depth, max_depth, sampling_invdepths,
                                      sampling_type, plane_range, debug, out, act=(float(fused_leaky_relu), key_activated_out))
        if self.warp_only and torch.is_grad_enabled() and any(f.requires_grad for f in feat_sources):
            return self._forward_warp_only_with_grad(feat_key, intrinsics_key, list(feat_sources), source_to_key_transforms, intrinsics_sources,
                                                     dict(num_sampling_points=num_sampling_points, min_depth=min_depth, max_depth=max_depth,
                                                          sampling_invdepths=sampling_invdepths, sampling_type=sampling_type,
                                                          plane_range=plane_range))
        if not needs_grad:
            return self._forward_impl(feat_key, intrinsics_key, feat_sources, source_to_key_transforms,
                                      intrinsics_sources, num_sampling_points, min_depth, max_depth, sampling_invdepths,
                                      sampling_type, plane_range, debug, out)
        assert out is None, "out= is not available on the differentiable path"
        mode = _corr_norm_mode(self.normalize)
        feat_sources = list(feat_sources)
        if mode == NORM_BEFORE:  # normalize(feat, dim=1) (:8-10, :170-172) through autograd, kernels see NORM_NONE
            feat_key = feat_key / (torch.linalg.norm(feat_key, dim=1, keepdim=True) + 1e-9)
            feat_sources = [f / (torch.linalg.norm(f, dim=1, keepdim=True) + 1e-9) for f in feat_sources]
            mode = NORM_NONE
        args = dict(intrinsics_key=intrinsics_key, source_to_key_transforms=source_to_key_transforms,
                    intrinsics_sources=intrinsics_sources, num_sampling_points=num_sampling_points, min_depth=min_depth,
                    max_depth=max_depth, sampling_invdepths=sampling_invdepths, sampling_type=sampling_type,
                    plane_range=plane_range, debug=debug)
        corr, mask = _CorrFunction.apply(self, args, mode, feat_key, *feat_sources)
        self.corrs = list(corr.unbind(0))
        self.masks = list(mask.unbind(0))
        return self.corrs, self.masks, self._inv

    def _forward_warp_only_with_grad(self, feat_key, intrinsics_key, feat_sources, source_to_key_transforms, intrinsics_sources, sampling):
        """``warp_only=True`` for inputs that require grad (WarpOnlyCorr.forward, :107-140): per view one un-normalised kernel call as an
        autograd node (rmvd_psweep_warp_fwd / rmvd_psweep_warp_bwd); the L2 normalisations run as torch ops around it -- "before" on the
        source map (:128-129), "after" / True on the warped map (:135-136; the kernel's output is already multiplied by the 0 / 1 mask,
        which the normalisation maps to itself)."""
        mode = _warp_norm_mode(self.normalize)
        args = dict(feat_key=feat_key.detach(), intrinsics_key=intrinsics_key, source_to_key_transforms=list(source_to_key_transforms),
                    intrinsics_sources=None if intrinsics_sources is None else list(intrinsics_sources), sampling=sampling)
        corrs, masks, inv, coeffs, points = [], [], None, [], []
        for v, f in enumerate(feat_sources):
            if mode == NORM_BEFORE:
                f = f / (torch.linalg.norm(f, dim=1, keepdim=True) + 1e-9)
            if f.requires_grad:
                warped, mask = _WarpFunction.apply(self, args, v, f)
                inv = self._inv_last
            else:
                c, m, inv = self._forward_impl(feat_key=args["feat_key"], feat_sources=[f], norm_mode=NORM_NONE, intrinsics_key=intrinsics_key,
                                               source_to_key_transforms=[args["source_to_key_transforms"][v]],
                                               intrinsics_sources=None if args["intrinsics_sources"] is None else [args["intrinsics_sources"][v]],
                                               **sampling)
                warped, mask = c[0], m[0]
            if mode == NORM_AFTER:
                warped = warped / (torch.linalg.norm(warped, dim=2, keepdim=True) + 1e-9)
            corrs.append(warped)
            masks.append(mask)
            coeffs += self.coeffs
            points += self.sampling_points
        self.corrs, self.masks, self.coeffs, self.sampling_points = corrs, masks, coeffs, points
        return self.corrs, self.masks, inv

    @torch.no_grad()
    def _warp_backward_impl(self, grad_warped, src_shape, key_shape, intrinsics_key, intrinsics_source, source_to_key_transform, inv, plane_range):
        dev = grad_warped.device
        N, C, hs, ws = src_shape
        h, w = key_shape[2], key_shape[3]
        inv_c = _f32c(inv)
        S = inv_c.shape[1]
        s0, s1 = (0, S) if plane_range is None else plane_range
        g = _f32c(grad_warped)
        assert g.shape == (N, s1 - s0, C, h, w)
        Kk = _calib(intrinsics_key, N, 3, dev, "intrinsics_key")
        Ks = _calib(intrinsics_source, N, 3, dev, "intrinsics_sources")
        T = _calib(source_to_key_transform, N, 4, dev, "source_to_key_transforms")
        grad = torch.empty((N, C, hs, ws), dtype=torch.float32, device=dev)
        st = _lib.load().rmvd_psweep_warp_bwd(_lib.ptr(Kk), _lib.ptr(Ks), _lib.ptr(T), _lib.ptr(inv_c), _lib.i64_array(_inv_strides(inv_c, N)),
                                              N, C, h, w, hs, ws, S, s0, s1, NORM_NONE, _lib.ptr(g), s1 - s0, 0, _lib.ptr(grad),
                                              _lib.stream_ptr())
        _lib.check(st, "rmvd_psweep_warp_bwd")
        return grad

    @torch.no_grad()
    def _backward_impl(self, grad_corr, feat_key, feat_sources, intrinsics_key, source_to_key_transforms,
                       intrinsics_sources, inv, plane_range, mode, need_key, need_src):
        """Gradients of ``sum <grad_corr, corr>`` with respect to ``feat_key`` and every ``feat_sources[v]``."""
        dev = feat_key.device
        if intrinsics_sources is None:
            intrinsics_sources = [intrinsics_key] * len(feat_sources)
        inv_c = _f32c(inv)
        feat_key = _f32c(feat_key)
        N, C, h, w = feat_key.shape
        S = inv_c.shape[1]
        V = len(feat_sources)
        s0, s1 = (0, S) if plane_range is None else plane_range
        Sl = s1 - s0
        srcs = [_f32c(f) for f in feat_sources]
        Kk = _calib(intrinsics_key, N, 3, dev, "intrinsics_key")
        Ks = [_calib(k, N, 3, dev, "intrinsics_sources") for k in intrinsics_sources]
        Ts = [_calib(t, N, 4, dev, "source_to_key_transforms") for t in source_to_key_transforms]
        hs = [f.shape[2] for f in srcs]
        ws = [f.shape[3] for f in srcs]
        grad_corr = _f32c(grad_corr)
        assert grad_corr.shape == (V, N, Sl, h, w)
        gkey = torch.empty_like(feat_key) if need_key else None
        gsrcs = [torch.empty_like(f) if need_src[v] else None for v, f in enumerate(srcs)]
        lib = _lib.load()
        nb = lib.rmvd_psweep_bwd_workspace_bytes(N, C, h, w, V, _lib.int_array(hs), _lib.int_array(ws))
        wsp = _workspace(nb, dev)
        st = lib.rmvd_psweep_corr_bwd(
            _lib.ptr(feat_key), _lib.ptr_array(srcs), _lib.ptr(Kk), _lib.ptr_array(Ks), _lib.ptr_array(Ts),
            _lib.ptr(inv_c), _lib.i64_array(_inv_strides(inv_c, N)), N, C, h, w, V, _lib.int_array(hs),
            _lib.int_array(ws), S, s0, s1, mode, _lib.ptr(grad_corr), Sl, 0, _lib.ptr(gkey),
            _lib.ptr_array_opt(gsrcs), _lib.ptr(wsp), wsp.numel(), _lib.stream_ptr())
        _lib.check(st, "rmvd_psweep_corr_bwd")
        return gkey, gsrcs

    @torch.no_grad()
    def _forward_impl(self, feat_key, intrinsics_key, feat_sources, source_to_key_transforms, intrinsics_sources=None,
                      num_sampling_points=None, min_depth=None, max_depth=None, sampling_invdepths=None,
                      sampling_type="linear_invdepth", plane_range=None, debug=False, out=None, norm_mode=None, peers=None, act=None):
        self.reset()
        dev = feat_key.device
        if intrinsics_sources is None:  # :441-442
            intrinsics_sources = [intrinsics_key] * len(feat_sources)
        assert len(feat_sources) == len(source_to_key_transforms) == len(intrinsics_sources)  # :444
        inv = self.get_sampling_invdepths(num_sampling_points, min_depth, max_depth, sampling_invdepths,
                                          sampling_type, device=dev)
        inv_c = _f32c(inv)
        feat_key = _f32c(feat_key)
        N, C, h, w = feat_key.shape
        S = inv_c.shape[1]
        assert inv_c.shape[2] in (1, h) and inv_c.shape[3] in (1, w)
        V = len(feat_sources)
        s0, s1 = (0, S) if plane_range is None else plane_range
        Sl = s1 - s0
        srcs = [_f32c(f) for f in feat_sources]
        Kk = _calib(intrinsics_key, N, 3, dev, "intrinsics_key")
        Ks = [_calib(k, N, 3, dev, "intrinsics_sources") for k in intrinsics_sources]
        Ts = [_calib(t, N, 4, dev, "source_to_key_transforms") for t in source_to_key_transforms]
        for f in srcs:
            if f.ndim != 4 or f.shape[0] != N or f.shape[1] != C:
                raise _lib.RmvdError(f"feat_sources: expected [{N},{C},hs,ws], got {tuple(f.shape)}")
        hs = [f.shape[2] for f in srcs]
        ws = [f.shape[3] for f in srcs]
        ist = _lib.i64_array(_inv_strides(inv_c, N))
        lib = _lib.load()
        lazy = [_LazyView((Kk, Ks[v], Ts[v], inv_c, N, h, w, hs[v], ws[v], S)) for v in range(V)]
        self.coeffs = [EpipolarCoeffs(z) for z in lazy]
        self.sampling_points = [EpipolarSamplingPoints(z) for z in lazy]
        if V == 0:
            if act is not None and act[1] is not None:
                act[1].copy_(torch.nn.functional.leaky_relu(feat_key, act[0]))
            return self.corrs, self.masks, inv
        if peers is not None:
            # depth-slab mode (parallel.SlabGather): rows go to every rank's [V,N,S,h,w] tensors from inside the kernel
            assert not self.warp_only and not debug and out is None
            nb = lib.rmvd_psweep_workspace_bytes(N, C, h, w, V, _lib.int_array(hs), _lib.int_array(ws))
            wsp = _workspace(nb, dev)
            G = len(peers["corr"])
            if peers.get("queue") is not None:       # push transport: a pusher kernel on peers["push_stream"] forwards the chunks
                q = peers["queue"]
                st = lib.rmvd_psweep_corr_fwd_push(
                    _lib.ptr(feat_key), _lib.ptr_array(srcs), _lib.ptr(Kk), _lib.ptr_array(Ks), _lib.ptr_array(Ts),
                    _lib.ptr(inv_c), ist, N, C, h, w, V, _lib.int_array(hs), _lib.int_array(ws), S, s0, s1,
                    _corr_norm_mode(self.normalize) if norm_mode is None else norm_mode, S, s0, G, peers["rank"],
                    _lib.raw_ptr_array(peers["corr"]), _lib.raw_ptr_array(peers["mask"]), _lib.raw_ptr_array(peers["flags"]),
                    peers["epoch"], _lib.ptr(q), q.numel() * q.element_size(), int(peers.get("push_ctas", 8)), _lib.ptr(wsp),
                    wsp.numel(), _lib.stream_ptr(), ctypes.c_void_p(peers["push_stream"].cuda_stream))
                _lib.check(st, "rmvd_psweep_corr_fwd_push")
                return None, None, inv
            st = lib.rmvd_psweep_corr_fwd_peers(
                _lib.ptr(feat_key), _lib.ptr_array(srcs), _lib.ptr(Kk), _lib.ptr_array(Ks), _lib.ptr_array(Ts),
                _lib.ptr(inv_c), ist, N, C, h, w, V, _lib.int_array(hs), _lib.int_array(ws), S, s0, s1,
                _corr_norm_mode(self.normalize) if norm_mode is None else norm_mode, S, s0, G, peers["rank"],
                _lib.raw_ptr_array(peers["corr"]), _lib.raw_ptr_array(peers["mask"]), _lib.raw_ptr_array(peers["flags"]),
                peers["epoch"], int(peers.get("partition", 0)), ctypes.c_void_p(peers.get("corr_mc") or None),
                ctypes.c_void_p(peers.get("mask_mc") or None), _lib.ptr(wsp), wsp.numel(), _lib.stream_ptr())
            _lib.check(st, "rmvd_psweep_corr_fwd_peers")
            return None, None, inv
        if not self.warp_only:
            out_S, out_s0 = Sl, 0
            if out is None:
                corr = torch.empty((V, N, Sl, h, w), dtype=torch.float32, device=dev)
                mask = torch.empty_like(corr)
            else:
                corr, mask = out
                if corr.shape[2] == S and Sl != S:      # full-height tensors: the slab lands at its own plane offset
                    out_S, out_s0 = S, s0
                for o in (corr, mask):
                    assert o.shape == (V, N, out_S, h, w) and o.dtype == torch.float32 and o.is_contiguous() and o.device == dev
            dbg = [None, None, None]
            if debug:
                dbg = [torch.empty((V, N, Sl, h, w), dtype=torch.int32, device=dev) for _ in range(2)]
                dbg.append(torch.empty((V, N, Sl, h, w), dtype=torch.uint8, device=dev))
            nb = lib.rmvd_psweep_workspace_bytes(N, C, h, w, V, _lib.int_array(hs), _lib.int_array(ws))
            wsp = _workspace(nb, dev)
            if act is not None:
                slope, key_out = act
                if key_out is not None:
                    assert key_out.shape == feat_key.shape and key_out.dtype == torch.float32 and key_out.is_contiguous() and \
                        key_out.device == dev and key_out.data_ptr() != feat_key.data_ptr()
                st = lib.rmvd_psweep_corr_fwd_act(
                    _lib.ptr(feat_key), _lib.ptr_array(srcs), slope, _lib.ptr(key_out), _lib.ptr(Kk), _lib.ptr_array(Ks),
                    _lib.ptr_array(Ts), _lib.ptr(inv_c), ist, N, C, h, w, V, _lib.int_array(hs), _lib.int_array(ws), S, s0, s1,
                    _corr_norm_mode(self.normalize) if norm_mode is None else norm_mode, _lib.ptr(corr), _lib.ptr(mask), out_S, out_s0,
                    _lib.ptr(wsp), wsp.numel(), _lib.stream_ptr())
                if st == -3:       # RMVD_E_MODE: shapes outside the persistent kernel -> activate first, then the ordinary call
                    feat_key = torch.nn.functional.leaky_relu(feat_key, slope)
                    srcs = [torch.nn.functional.leaky_relu(f, slope) for f in srcs]
                    if key_out is not None:
                        key_out.copy_(feat_key)
                    act = None
                else:
                    _lib.check(st, "rmvd_psweep_corr_fwd_act")
            st = 0 if act is not None else lib.rmvd_psweep_corr_fwd(
                _lib.ptr(feat_key), _lib.ptr_array(srcs), _lib.ptr(Kk), _lib.ptr_array(Ks), _lib.ptr_array(Ts),
                _lib.ptr(inv_c), ist, N, C, h, w, V, _lib.int_array(hs), _lib.int_array(ws), S, s0, s1,
                _corr_norm_mode(self.normalize) if norm_mode is None else norm_mode, _lib.ptr(corr), _lib.ptr(mask), out_S, out_s0,
                _lib.ptr(dbg[0]),
                _lib.ptr(dbg[1]), _lib.ptr(dbg[2]), _lib.ptr(wsp), wsp.numel(), _lib.stream_ptr())
            _lib.check(st, "rmvd_psweep_corr_fwd")
            self.corrs = list(corr.unbind(0))
            self.masks = list(mask.unbind(0))
            self._stacked = (corr, mask)
            if debug:
                self.debug = dict(ix0=dbg[0], iy0=dbg[1], geom=dbg[2])
        else:
            for v in range(V):
                warped = torch.empty((N, Sl, C, h, w), dtype=torch.float32, device=dev)
                mask = torch.empty((N, Sl, h, w), dtype=torch.float32, device=dev)
                st = lib.rmvd_psweep_warp_fwd(
                    _lib.ptr(srcs[v]), _lib.ptr(Kk), _lib.ptr(Ks[v]), _lib.ptr(Ts[v]), _lib.ptr(inv_c), ist, N, C, h,
                    w, hs[v], ws[v], S, s0, s1, _warp_norm_mode(self.normalize) if norm_mode is None else norm_mode, _lib.ptr(warped),
                    _lib.ptr(mask), Sl, 0, _lib.stream_ptr())
                _lib.check(st, "rmvd_psweep_warp_fwd")
                self.corrs.append(warped)
                self.masks.append(mask)
        return self.corrs, self.masks, inv
