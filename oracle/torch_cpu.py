"""Oracle, multi-threaded CPU variant (TEST / BASELINE INFRASTRUCTURE ONLY -- never imported by the product).

The reference is Python on top of ATen; its own files cannot travel to the GPU box (``/root/reference`` is absent
there), so the timed CPU baseline of ``bench.py`` is this restatement of the SAME operator sequence on torch CPU
tensors with all host threads: all-pairs ``matmul`` then ``grid_sample`` of the correlation matrix for robust_mvd
(planesweep_corr.py:152-195, :85-102, :333-349, :489-506) and ``homo_warp`` + running sum / sum of squares for
MVSNet (blocks/utils.py:222-268, mvsnet.py:123-160).  Pinned against the numpy oracle and the golden vectors in
``tests/test_oracle_golden.py``.  Geometry (twelve scalars per view) comes from the numpy oracle.
"""
import numpy as np
import torch
import torch.nn.functional as F

from . import planesweep as _ps


def _grid_sample_with_mask(x, us, vs):
    """warp() grid branch, zeros padding (planesweep_corr.py:85-102).  x [K,1,hs,ws]; us, vs [K,S,1]."""
    hs, ws = x.shape[-2:]
    grid = torch.stack((2 * us / ws - 1, 2 * vs / hs - 1), -1)  # K,S,1,2
    out = F.grid_sample(x, grid, padding_mode="zeros", align_corners=False)
    ones = torch.ones((x.shape[0], 1, hs, ws), device=x.device)
    mask = F.grid_sample(ones, grid, padding_mode="zeros", align_corners=False)
    mask[mask < 0.9999] = 0
    mask[mask > 0] = 1
    return out, mask


def planesweep_correlation(feat_key, intrinsics_key, feat_sources, source_to_key_transforms, intrinsics_sources=None,
                           num_sampling_points=None, min_depth=None, max_depth=None, normalize="dim", device=None,
                           return_torch=False):
    """PlanesweepCorrelation.forward (warp_only=False) with the reference's algorithm on stock torch ops.  numpy in; numpy out
    (default) or torch tensors left on ``device`` (``return_torch=True``).  ``device=None``: CPU, all host threads -- the timed
    CPU baseline.  ``device="cuda"``: the same operator sequence on the GPU -- the "stock PyTorch on B200" baseline of SURVEY.md
    section 8(d) and the same-device oracle for the masks (CUDA-ATen grid_sample, section 7.3-1)."""
    dev = torch.device("cpu" if device is None else device)
    fk = torch.as_tensor(feat_key) if isinstance(feat_key, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(feat_key))
    fk = fk.to(dev)
    N, C, h, w = fk.shape
    if intrinsics_sources is None:
        intrinsics_sources = [intrinsics_key] * len(feat_sources)
    inv = _ps.compute_sampling_invdepths(min_depth, max_depth, num_sampling_points)
    ds = torch.from_numpy(inv)[:, :, None, None].to(dev)
    corrs, masks = [], []
    for v, T in enumerate(source_to_key_transforms):
        fs = feat_sources[v] if isinstance(feat_sources[v], torch.Tensor) else torch.from_numpy(np.ascontiguousarray(feat_sources[v]))
        fs = fs.to(dev)
        hs, ws = fs.shape[-2:]
        co = [torch.from_numpy(c).to(dev) for c in _ps.epipolar_coeffs(intrinsics_key, intrinsics_sources[v], T, h, w, hs, ws)]
        u_h, v_h, k_h, m_u, m_v, m_k = co
        den = k_h + m_k * ds
        us = (u_h + m_u * ds) / den
        vs = (v_h + m_v * ds) / den
        for t in (us, vs):  # :336-338
            t[torch.isinf(t)] = 1e9 * torch.sign(t)[torch.isinf(t)]
            t[torch.isnan(t)] = 1e9
        zs = (1.0 / ds) * torch.ones_like(us)
        z_poles = -(m_k / k_h)
        geom = (zs > 0) & (((k_h > 0) & (zs > z_poles)) | ((k_h < 0) & (zs < z_poles)) | ((k_h == 0) & (m_k > 0)))
        a = fk.view(N, C, h * w)
        b = fs.view(N, C, hs * ws)
        if normalize and normalize != "dim":
            a = a / (torch.linalg.norm(a, dim=1, keepdim=True) + 1e-9)
            b = b / (torch.linalg.norm(b, dim=1, keepdim=True) + 1e-9)
        corr = torch.matmul(a.transpose(1, 2), b)  # :174
        if normalize == "dim":
            corr = corr / torch.sqrt(torch.tensor(C, device=dev).float())
        corr = corr.reshape(N * h * w, 1, hs, ws)
        S = us.shape[1]
        gu = us.permute(0, 2, 3, 1).reshape(N * h * w, S, 1)
        gv = vs.permute(0, 2, 3, 1).reshape(N * h * w, S, 1)
        c, m = _grid_sample_with_mask(corr, gu, gv)
        c = c.reshape(N, h, w, S).permute(0, 3, 1, 2)
        m = m.reshape(N, h, w, S).permute(0, 3, 1, 2)
        c = c * m
        c = c * geom
        m = m * geom
        corrs.append(c.contiguous() if return_torch else c.contiguous().cpu().numpy())
        masks.append(m.contiguous() if return_torch else m.contiguous().cpu().numpy())
    return corrs, masks, inv


def mvs_variance(feats, proj_mats, depth_values, device=None, return_torch=False):
    """mvsnet.py:123-136 + homo_warp (utils.py:222-268) on stock torch ops (CPU by default, see planesweep_correlation)."""
    dev = torch.device("cpu" if device is None else device)
    as_t = lambda a: a.to(dev) if torch.is_tensor(a) else torch.from_numpy(np.ascontiguousarray(a)).to(dev)  # noqa: E731  (tensors keep their graph)
    feats, proj, depth = as_t(feats), as_t(proj_mats), as_t(depth_values)
    B, V, C, H, W = feats.shape
    D = depth.shape[1]
    ys, xs = torch.meshgrid(torch.arange(H, dtype=torch.float32, device=dev), torch.arange(W, dtype=torch.float32, device=dev),
                            indexing="ij")
    ref_grid = torch.stack((xs.reshape(-1), ys.reshape(-1), torch.ones(H * W, device=dev)), 0)[None].expand(B, -1, -1)
    vol = feats[:, 0].unsqueeze(2).repeat(1, 1, D, 1, 1)
    vsum, vsq = vol, vol ** 2
    for v in range(1, V):
        tr = proj[:, v] @ proj[:, 0]
        R, T = tr[:, :3, :3], tr[:, :3, 3:]
        g = (ref_grid.unsqueeze(2) * depth.view(B, 1, D, 1)).view(B, 3, D * H * W)
        sg = R @ g + T
        sg = sg[:, :2] / sg[:, -1:]
        sg[:, 0] = sg[:, 0] / ((W - 1) / 2) - 1
        sg[:, 1] = sg[:, 1] / ((H - 1) / 2) - 1
        sg = sg.permute(0, 2, 1).view(B, D, H * W, 2)
        wv = F.grid_sample(feats[:, v], sg, mode="bilinear", padding_mode="zeros", align_corners=False).view(B, C, D, H, W)
        vsum = vsum + wv
        vsq = vsq + wv ** 2
    out = vsq.div_(V).sub_(vsum.div_(V).pow_(2))
    return out if return_torch else out.cpu().numpy()


def mvs_regress(cost, depth_values, device=None, return_torch=False):
    """softmax + depth_regression + 4-bin confidence (mvsnet.py:140-160, utils.py:271-274) on stock torch ops.  numpy in/out
    (tensors in / out with return_torch)."""
    dev = torch.device("cpu" if device is None else device)
    as_t = lambda a: a.to(dev) if torch.is_tensor(a) else torch.from_numpy(np.ascontiguousarray(a)).to(dev)  # noqa: E731
    cost, depth = as_t(cost), as_t(depth_values)
    prob = F.softmax(cost, dim=1)
    d = torch.sum(prob * depth.view(*depth.shape, 1, 1), 1)
    D = cost.shape[1]
    p4 = 4 * F.avg_pool3d(F.pad(prob.unsqueeze(1), pad=(0, 0, 0, 0, 1, 2)), (4, 1, 1), stride=1, padding=0).squeeze(1)
    idx = torch.sum(prob * torch.arange(D, device=dev, dtype=torch.float32).view(1, D, 1, 1), 1).long()
    conf = torch.gather(p4, 1, idx.unsqueeze(1)).squeeze(1)
    if return_torch:
        return d, conf
    return d.cpu().numpy(), conf.cpu().numpy()
