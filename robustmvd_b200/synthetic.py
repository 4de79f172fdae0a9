"""Seeded synthetic inputs of the shapes named in BASELINE.json.

Used by ``bench.py``, the tests and the golden-vector tooling, so that every party sees identical inputs.
Pure numpy: importable without a GPU and without the CUDA library.
"""
import numpy as np


def rot_y(deg):
    a = np.deg2rad(deg)
    c, s = np.cos(a), np.sin(a)
    return np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]], dtype=np.float64)


def rot_x(deg):
    a = np.deg2rad(deg)
    c, s = np.cos(a), np.sin(a)
    return np.array([[1, 0, 0], [0, c, -s], [0, s, c]], dtype=np.float64)


def rot_z(deg):
    a = np.deg2rad(deg)
    c, s = np.cos(a), np.sin(a)
    return np.array([[c, -s, 0], [s, c, 0], [0, 0, 1]], dtype=np.float64)


def make_transform(R=None, t=(0.0, 0.0, 0.0), n=1):
    """4x4 key->source transform ``[n,4,4]`` float32 (rmvd/data/README.md: p_view = T @ p_key)."""
    T = np.eye(4, dtype=np.float64)
    if R is not None:
        T[:3, :3] = R
    T[:3, 3] = t
    return np.repeat(T[None].astype(np.float32), n, 0)


def relative_intrinsics(h, w, fx_rel=0.58, fy_rel=None, cx_rel=0.5, cy_rel=0.5, n=1):
    """Relative intrinsics as robust_mvd's input adapter produces them (robust_mvd.py:117-120)."""
    if fy_rel is None:
        fy_rel = fx_rel * w / h
    K = np.array([[fx_rel, 0, cx_rel], [0, fy_rel, cy_rel], [0, 0, 1]], dtype=np.float32)
    return np.repeat(K[None], n, 0)


def kitti_like_views(num_views, n=1):
    """Source poses of SURVEY.md 8(d): small yaw (<= 5 deg) + x-translation of +-0.54 m*k + 0.05 m*k forward."""
    out = []
    for k in range(1, num_views + 1):
        sign = -1.0 if (k % 2) else 1.0
        yaw = sign * min(5.0, 0.6 * k)
        out.append(make_transform(rot_y(yaw), (sign * 0.54 * ((k + 1) // 2), 0.0, 0.05 * k), n))
    return out


def features(shape, seed, relu=False):
    rng = np.random.default_rng(seed)
    x = rng.standard_normal(shape, dtype=np.float32)
    if relu:
        x = np.maximum(x, 0)
    return x


def planesweep_inputs(h, w, C, V, n=1, seed=1234, hs=None, ws=None, relu=False):
    """Inputs of ``PlanesweepCorrelation.forward`` at a given feature shape (cfg2: h,w=48,160 C=256 V=2)."""
    hs = h if hs is None else hs
    ws = w if ws is None else ws
    return dict(
        feat_key=features((n, C, h, w), seed, relu),
        intrinsics_key=relative_intrinsics(h, w, n=n),
        feat_sources=[features((n, C, hs, ws), seed + 1 + v, relu) for v in range(V)],
        transforms=kitti_like_views(V, n),
        intrinsics_sources=None,
    )


CFG2 = dict(h=48, w=160, C=256, V=2, S=256, min_depth=0.4, max_depth=1000.0)   # KITTI 384x1280 / 8
CFG4 = dict(h=96, w=144, C=256, V=2, S=256, min_depth=0.4, max_depth=1000.0)   # ETH3D 768x1152 / 8
CFG3 = dict(h=224, w=304, C=32, V=5, D=192, min_depth=0.425, max_depth=0.935)  # DTU 896x1216 / 4


def mvs_inputs(h, w, C, V, D, n=1, seed=4321, min_depth=0.425, max_depth=0.935, fx=360.0):
    """Inputs of the MVSNet volume (mvsnet.py:109-135): feats ``[B,V,C,h,w]`` (view 0 = key), projection
    matrices ``[B,V,4,4]`` at feature resolution with the key view already inverted (:82-86), depth samples
    ``[B,D]`` linear in depth (:66-71).  DTU-like: fx = 360 px at 1/4 res, baselines 0.05-0.2 m."""
    rng = np.random.default_rng(seed)
    feats = rng.standard_normal((n, V, C, h, w), dtype=np.float32)
    K = np.array([[fx, 0, w / 2.0, 0], [0, fx, h / 2.0, 0], [0, 0, 1, 0], [0, 0, 0, 1]], dtype=np.float64)
    proj = np.zeros((n, V, 4, 4), np.float32)
    for v in range(V):
        if v == 0:
            pose = np.eye(4)
        else:
            sign = -1.0 if (v % 2) else 1.0
            pose = np.eye(4)
            pose[:3, :3] = rot_y(sign * 2.0 * ((v + 1) // 2)) @ rot_x(0.7 * v)
            pose[:3, 3] = (sign * 0.05 * (1 + v // 2), 0.02 * v, 0.005 * v)
        P = pose.astype(np.float32).copy()
        P[:3, :4] = (K[:3, :3].astype(np.float32) @ P[:3, :4]).astype(np.float32)
        if v == 0:
            P = np.linalg.inv(P.astype(np.float64)).astype(np.float32)
        proj[:, v] = P
    depth = np.repeat(np.linspace(min_depth, max_depth, D, dtype=np.float32)[None], n, 0)
    cost = rng.standard_normal((n, D, h, w), dtype=np.float32) * np.float32(3.0)
    return dict(feats=feats, proj=proj, depth=depth, cost=cost)


def model_sample(H, W, num_sources, seed=2024, batch=None, depth_range=None):
    """A synthetic ``model.run`` sample in the reference's data format (rmvd/data/README.md): list of ``3xHxW`` float32
    images in [0,255] (key view first), ``3x3`` pixel intrinsics, ``4x4`` key->view poses, ``keyview_idx = 0``.
    Smooth random textures so the encoders see image-like input; KITTI-like intrinsics and baselines.
    ``batch=None`` -> un-batched sample; ``batch=n`` -> every entry gets a leading dimension n."""
    rng = np.random.default_rng(seed)
    n = 1 if batch is None else batch
    ys, xs = np.meshgrid(np.linspace(0, 1, H, dtype=np.float32), np.linspace(0, 1, W, dtype=np.float32), indexing="ij")
    images = []
    for v in range(num_sources + 1):
        im = np.zeros((n, 3, H, W), np.float32)
        for k in range(6):
            fx, fy, ph = rng.uniform(1, 9, (n, 3, 1, 1)), rng.uniform(1, 9, (n, 3, 1, 1)), rng.uniform(0, 6.28, (n, 3, 1, 1))
            im += (np.sin(6.28 * (fx * xs + fy * ys) + ph) / (k + 1)).astype(np.float32)
        im = 127.5 + 50.0 * im + rng.normal(0, 4.0, im.shape).astype(np.float32)
        images.append(np.clip(im, 0, 255).astype(np.float32))
    K = np.array([[0.58 * W, 0, 0.5 * W], [0, 0.58 * W, 0.5 * H], [0, 0, 1]], np.float32)
    intrinsics = [np.repeat(K[None], n, 0) for _ in range(num_sources + 1)]
    poses = [make_transform(n=n)] + kitti_like_views(num_sources, n)
    sample = dict(images=images, intrinsics=intrinsics, poses=poses, keyview_idx=0 if batch is None else np.zeros(n, np.int64))
    if depth_range is not None:
        sample["depth_range"] = tuple(np.float32(d) if batch is None else np.full(n, d, np.float32) for d in depth_range)
    if batch is None:
        for key in ("images", "intrinsics", "poses"):
            sample[key] = [a[0] for a in sample[key]]
    return sample
