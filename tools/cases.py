"""Seeded parity cases (shapes x poses x options) shared by the oracle validation, the golden-vector
generator and the GPU parity tests.  Pure numpy."""
import numpy as np

from robustmvd_b200 import synthetic as syn


def _ps(name, h, w, C, transforms, S=16, seed=0, hs=None, ws=None, n=1, K_key=None, K_src=None,
        kwargs=None, **opts):
    hs = h if hs is None else hs
    ws = w if ws is None else ws
    V = len(transforms)
    c = dict(
        name=name,
        feat_key=syn.features((n, C, h, w), seed),
        intrinsics_key=syn.relative_intrinsics(h, w, n=n) if K_key is None else K_key,
        feat_sources=[syn.features((n, C, hs, ws), seed + 1 + v) for v in range(V)],
        transforms=transforms,
        intrinsics_sources=K_src,
        kwargs=kwargs if kwargs is not None else dict(num_sampling_points=S, min_depth=0.4, max_depth=1000.0),
    )
    c.update(opts)
    return c


def planesweep_cases(include_large=True):
    T = syn.make_transform
    cases = []
    # degenerate poses
    cases.append(_ps("identity_6x10", 6, 10, 8, [T()], seed=1))
    cases.append(_ps("stereo_x_12x20", 12, 20, 8, [T(t=(-0.54, 0, 0)), T(t=(0.54, 0, 0))], seed=2))
    cases.append(_ps("forward_12x20", 12, 20, 8, [T(t=(0, 0, 0.5)), T(t=(0, 0, -0.5))], seed=3))
    cases.append(_ps("rot90_fw_12x20", 12, 20, 8, [T(syn.rot_y(90), (0.3, 0, 0.2))], seed=4))
    cases.append(_ps("behind_12x20", 12, 20, 8, [T(syn.rot_y(180), (0, 0, 1.0)), T(t=(0, 0, -3.0))], seed=5))
    cases.append(_ps("roll_pitch_12x20", 12, 20, 8,
                     [T(syn.rot_z(12) @ syn.rot_x(6), (0.2, -0.1, 0.03))], seed=6))
    # non-square intrinsics, off-centre principal point, source resolution != key resolution, own source K
    Kk = syn.relative_intrinsics(12, 20, fx_rel=0.7, fy_rel=1.1, cx_rel=0.45, cy_rel=0.55)
    Ks = syn.relative_intrinsics(9, 14, fx_rel=0.62, fy_rel=1.0, cx_rel=0.52, cy_rel=0.48)
    cases.append(_ps("src_res_differs", 12, 20, 8, [T(syn.rot_y(3), (0.3, 0.02, 0.01)), T(syn.rot_y(-4), (-0.4, 0, 0.05))],
                     hs=9, ws=14, K_key=Kk, K_src=[Ks, Ks], seed=7))
    # batch of 2 with per-sample depth ranges (compute_sampling_invdepths with arrays)
    cases.append(_ps("batch2_ranges", 12, 20, 8,
                     [T(syn.rot_y(2), (0.3, 0, 0), n=2), T(syn.rot_y(-2), (-0.3, 0.05, 0.02), n=2)], n=2, seed=8,
                     kwargs=dict(num_sampling_points=12, min_depth=np.array([0.5, 1.0], np.float32),
                                 max_depth=np.array([50.0, 200.0], np.float32))))
    cases.append(_ps("linear_depth", 12, 20, 8, [T(t=(0.3, 0, 0))], seed=9,
                     kwargs=dict(num_sampling_points=10, min_depth=0.5, max_depth=20.0, sampling_type="linear_depth")))
    # explicit inverse depths of rank 2, 3, 4
    rng = np.random.default_rng(10)
    for rank, shape in ((2, (1, 9)), (3, (1, 9, 12)), (4, (1, 9, 12, 20))):
        inv = (rng.random(shape, dtype=np.float32) * np.float32(2.0) + np.float32(0.01))
        cases.append(_ps(f"explicit_inv_rank{rank}", 12, 20, 8, [T(syn.rot_y(2), (0.3, 0, 0.02))], seed=10 + rank,
                         kwargs=dict(sampling_invdepths=inv)))
    # zero and negative inverse depths (zs = inf / negative: visible_in_key false)
    inv = np.array([[0.0, -0.5, 1e-12, 0.3, 3.0]], np.float32)
    cases.append(_ps("inv_zero_negative", 12, 20, 8, [T(t=(0.3, 0, 0.1))], seed=15, kwargs=dict(sampling_invdepths=inv)))
    # normalisation variants
    for nm in (False, True, "before"):
        cases.append(_ps(f"normalize_{nm}", 12, 20, 8, [T(syn.rot_y(2), (0.3, 0, 0.02))], seed=20, normalize=nm))
    # warp only
    for nm in (False, "before", "after", "dim"):
        cases.append(_ps(f"warp_only_{nm}", 12, 20, 8, [T(syn.rot_y(2), (0.3, 0, 0.02)), T(t=(-0.3, 0, 0))], seed=21,
                         warp_only=True, normalize=nm))
    # C not a multiple of 4, odd sizes
    cases.append(_ps("odd_sizes", 7, 11, 5, [T(syn.rot_y(1), (0.2, 0, 0))], S=7, seed=22))
    # banded-GEMM kernel (C % 32 == 0) over the same pose zoo, plus geometries that overflow its band capacity
    cases.append(_ps("band_stereo_c32", 24, 40, 32, [T(t=(-0.54, 0, 0)), T(t=(0.54, 0, 0))], S=64, seed=50))
    cases.append(_ps("band_forward_c32", 24, 40, 32, [T(t=(0, 0, 0.5)), T(t=(0.02, 0.01, -0.5))], S=64, seed=51))
    cases.append(_ps("band_roll_c64", 20, 36, 64, [T(syn.rot_z(25) @ syn.rot_x(6), (0.3, -0.1, 0.03))], S=48, seed=52))
    cases.append(_ps("band_rot90_behind_c32", 16, 24, 32,
                     [T(syn.rot_y(90), (0.3, 0, 0.2)), T(syn.rot_y(180), (0, 0, 1.0)), T(t=(0, 0, -3.0))], S=24, seed=53))
    Kk2 = syn.relative_intrinsics(24, 40, fx_rel=0.7, fy_rel=1.1, cx_rel=0.45, cy_rel=0.55, n=2)
    Ks2 = syn.relative_intrinsics(18, 28, fx_rel=0.62, fy_rel=1.0, cx_rel=0.52, cy_rel=0.48, n=2)
    cases.append(_ps("band_srcres_batch2_c32", 24, 40, 32,
                     [T(syn.rot_y(3), (0.3, 0.02, 0.01), n=2), T(syn.rot_y(-4), (-0.4, 0, 0.05), n=2)], hs=18, ws=28,
                     K_key=Kk2, K_src=[Ks2, Ks2], n=2, seed=54,
                     kwargs=dict(num_sampling_points=40, min_depth=np.array([0.5, 1.0], np.float32),
                                 max_depth=np.array([50.0, 200.0], np.float32))))
    rng = np.random.default_rng(55)
    inv4 = (rng.random((1, 20, 24, 40), dtype=np.float32) * np.float32(2.0) + np.float32(0.01))
    cases.append(_ps("band_perpixel_inv_direct_c32", 24, 40, 32, [T(syn.rot_y(2), (2.5, 0, 0.02))], seed=56,
                     kwargs=dict(sampling_invdepths=inv4)))
    inv5 = (np.random.default_rng(60).random((1, 16, 24, 160), dtype=np.float32) * np.float32(1.2) + np.float32(0.01))
    cases.append(_ps("band_scatter_direct_c32", 24, 160, 32, [T(t=(1.0, 0.1, 0.0))], seed=60,
                     kwargs=dict(sampling_invdepths=inv5)))
    cases.append(_ps("band_split_forward_c32", 48, 96, 32, [T(syn.rot_z(8), (0.1, 0.05, 1.2))], S=256, seed=57))
    cases.append(_ps("band_norm_before_c32", 24, 40, 32, [T(syn.rot_y(2), (0.3, 0, 0.02))], S=32, seed=58, normalize=True))
    cases.append(_ps("band_ragged_tile_c32", 7, 11, 32, [T(syn.rot_y(1), (0.2, 0, 0))], S=37, seed=59))
    if include_large:
        # KITTI-like geometry at reduced channel count, full plane count
        cases.append(_ps("kitti_like_c32", 48, 160, 32, syn.kitti_like_views(2), S=256, seed=30))
        cases.append(_ps("eth3d_like_c16", 96, 144, 16, syn.kitti_like_views(3), S=64, seed=31))
    return cases


BACKWARD_CASE_NAMES = ("identity_6x10", "stereo_x_12x20", "forward_12x20", "behind_12x20", "roll_pitch_12x20",
                       "src_res_differs", "batch2_ranges", "linear_depth", "explicit_inv_rank4", "inv_zero_negative",
                       "normalize_False", "normalize_True", "normalize_before", "odd_sizes", "band_stereo_c32",
                       "band_roll_c64", "band_rot90_behind_c32", "band_srcres_batch2_c32")


def backward_cases():
    """Cases of the correlation backward: a subset of the forward zoo plus channel counts that take
    the vectorised kernel (C = 128 k).  ``grad_corrs(case)`` gives the seeded upstream gradients."""
    T = syn.make_transform
    by_name = {c["name"]: c for c in planesweep_cases(include_large=False)}
    out = [by_name[n] for n in BACKWARD_CASE_NAMES]
    out.append(_ps("bwd_c128_stereo", 12, 20, 128, [T(t=(-0.54, 0, 0)), T(syn.rot_y(2), (0.54, 0, 0.03))], S=24, seed=70))
    out.append(_ps("bwd_c256_kitti_small", 12, 40, 256, syn.kitti_like_views(2), S=48, seed=71))
    out.append(_ps("bwd_c128_batch2_srcres", 10, 18, 128, [T(syn.rot_z(5), (0.3, 0.02, 0.01), n=2)], hs=8, ws=13, n=2, S=20,
                   seed=72))
    # source map 12x finer than the key map: one plane of a 32-pixel tile spans > 320 texels (pixel-split path of the tiled
    # kernel); and a strongly rolled view whose band boxes only fit for a few planes (plane-halving path)
    out.append(_ps("bwd_zoom_split", 6, 32, 16, [T(syn.rot_y(1), (0.2, 0.01, 0.0))], hs=40, ws=400, S=10, seed=73))
    out.append(_ps("bwd_roll40_c32", 16, 64, 32, [T(syn.rot_z(40) @ syn.rot_x(4), (0.4, -0.1, 0.05))], S=64, seed=74))
    return out


def grad_corrs(case, S):
    """Seeded upstream gradients, one ``[N,S,h,w]`` array per source view."""
    import zlib
    n, _, h, w = case["feat_key"].shape
    rng = np.random.default_rng(zlib.crc32(case["name"].encode()))
    return [rng.standard_normal((n, S, h, w)).astype(np.float32) for _ in case["feat_sources"]]


def border_case():
    """Caller grids for TorchCorr(padding_mode="border"): samples up to 4 texels outside the 9 x 14 source map on every side, a few
    exactly on the rim, and a geometric mask."""
    rng = np.random.default_rng(808)
    N, C, h, w, hs, ws, S = 2, 16, 6, 10, 9, 14, 7
    us = rng.uniform(-4, ws + 4, (N, S, h, w)).astype(np.float32)
    vs = rng.uniform(-4, hs + 4, (N, S, h, w)).astype(np.float32)
    us[0, 0, 0, :4] = (0.0, 0.5, ws - 0.5, float(ws))
    vs[0, 0, 0, :4] = (0.0, 0.5, hs - 0.5, float(hs))
    return dict(feat_key=syn.features((N, C, h, w), 809), feat_src=syn.features((N, C, hs, ws), 810),
                grids=np.stack([us, vs], 2).astype(np.float32), geom=rng.random((N, S, h, w)) < 0.8)


def fullsize_planesweep_case():
    """BASELINE.json configs[3] at full size: ETH3D 768x1152 -> 96x144 features, C=256, 256 planes, 2 source views."""
    c = syn.CFG4
    return _ps("cfg4_eth3d_full", c["h"], c["w"], c["C"], syn.kitti_like_views(c["V"]), S=c["S"], seed=4004)


def fullsize_mvs_case():
    """BASELINE.json configs[2] at full size: DTU 896x1216 -> 224x304 features, C=32, 5 views, 192 planes."""
    c = syn.CFG3
    return dict(name="cfg3_dtu_full", **syn.mvs_inputs(c["h"], c["w"], c["C"], c["V"], c["D"]))


def mvs_cases(include_large=True):
    cases = []
    cases.append(dict(name="mvs_small", **syn.mvs_inputs(12, 16, 8, 3, 8, seed=40, fx=20.0, min_depth=0.5, max_depth=1.5)))
    cases.append(dict(name="mvs_batch2", **syn.mvs_inputs(10, 14, 4, 4, 6, n=2, seed=41, fx=18.0)))
    cases.append(dict(name="mvs_odd", **syn.mvs_inputs(7, 9, 5, 2, 5, seed=42, fx=10.0)))
    if include_large:
        cases.append(dict(name="mvs_dtu_quarter", **syn.mvs_inputs(56, 76, 32, 5, 48, seed=43, fx=90.0)))
    return cases


def warp_backward_cases():
    """Warp-only block (WarpOnlyCorr) under autograd: tools/make_warp_bwd_golden.py, tests/test_gpu_planesweep_bwd.py."""
    out = []
    for name, normalize, seed, (h, w, C, V, n) in (("plain", False, 71, (10, 14, 6, 2, 2)), ("before", "before", 72, (9, 13, 5, 2, 1)),
                                                   ("after", "after", 73, (8, 12, 7, 1, 2)), ("dim_c32", "dim", 74, (12, 20, 32, 2, 1))):
        inp = syn.planesweep_inputs(h, w, C, V, n=n, seed=seed)
        out.append(dict(name=name, normalize=normalize, feat_key=inp["feat_key"], feat_sources=inp["feat_sources"],
                        intrinsics_key=inp["intrinsics_key"], transforms=inp["transforms"],
                        kwargs=dict(num_sampling_points=12, min_depth=0.4, max_depth=1000.0)))
    return out


def warp_upstream_grads(case, shape):
    rng = np.random.default_rng(5000 + int(np.prod(shape)) % 9973)
    return [rng.standard_normal(shape, dtype=np.float32) for _ in case["feat_sources"]]


def mvs_upstream_grads(case):
    """Upstream gradients of the MVSNet backward cases (tools/make_mvs_bwd_golden.py, tests/test_gpu_mvs.py): for the variance
    volume ``[B,C,D,h,w]`` (a tenth of the entries exactly zero) and the regressed depth ``[B,h,w]``, from the case's shapes."""
    B, V, C, h, w = case["feats"].shape
    D = case["depth"].shape[1]
    rng = np.random.default_rng(1000 + B * 131 + C * 17 + D)
    g_var = rng.standard_normal((B, C, D, h, w), dtype=np.float32)
    g_var[rng.random(g_var.shape) < 0.1] = 0.0
    g_depth = rng.standard_normal((B, h, w), dtype=np.float32)
    return g_var, g_depth


# ---------------------------------------------------------------------------------------------------------
# model-level cases (tools/make_model_golden.py, tests/test_gpu_models.py)
# ---------------------------------------------------------------------------------------------------------
MODEL_SEED = 1234            # torch.manual_seed before create_model: both sides draw identical weights
MVSNET_PROB_GAIN = 3.0e5     # cost_regularization.prob.weight *= gain after the seeded init: with untrained weights the
                             # regularised cost is ~1e-5 and the softmax flat; the gain makes depth / confidence depend
                             # on the cost volume (std of depth ~ 5 m over a 2..40 m range)
MODEL_CASES = {
    "robust_mvd": dict(model="robust_mvd", sample=dict(H=128, W=192, num_sources=2, seed=11)),
    "robust_mvd_batch2": dict(model="robust_mvd", sample=dict(H=64, W=128, num_sources=3, seed=12, batch=2)),
    "mvsnet_train": dict(model="mvsnet_train", sample=dict(H=64, W=96, num_sources=2, seed=13, depth_range=(2.0, 40.0))),
    "mvsnet_train_default_range": dict(model="mvsnet_train", sample=dict(H=64, W=96, num_sources=3, seed=14)),
}


MODEL_GRAD_CASES = ("robust_mvd_batch2",)
MODEL_GRAD_PARAMS = ("encoder.conv1.0.weight", "encoder.conv3.0.weight", "encoder.conv3.0.bias",
                     "fusion_block.corr_to_view_weight.0.weight", "fusion_enc_block.conv3_1.0.weight",
                     "context_encoder.conv_redir.0.weight", "decoder.pred_5.0.weight")
MODEL_GRAD_STRIDE = 7


def model_grad_loss(aux):
    """Scalar the model-gradient goldens differentiate (same expression on both sides)."""
    return aux["invdepth"].mean() + 0.1 * aux["invdepth_log_b"].mean()


def prepare_model_for_case(model, model_name):
    """Test-only weight tweak applied identically to the reference model and to the candidate."""
    import torch
    if model_name == "mvsnet_train":
        with torch.no_grad():
            model.cost_regularization.prob.weight *= MVSNET_PROB_GAIN
    return model


def _vis_cam(R, t, fx, fy, cx, cy, n):
    cam = np.zeros((n, 2, 4, 4), np.float32)
    cam[:, 0] = np.eye(4, dtype=np.float32)
    cam[:, 0, :3, :3] = R
    cam[:, 0, :3, 3] = t
    cam[:, 1, :3, :3] = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1]], np.float32)
    cam[:, 1, 3, 3] = 1.0
    return cam


def vis_cases():
    """Vis-MVSNet cost-volume cases: reference / source features ``[n,c,h,w]``, cameras ``[n,2,4,4]``
    (``[:,0]`` world-to-camera extrinsics, ``[:,1]`` intrinsics), plane schedule ``depth_start + k * depth_interval``."""
    out = []

    def case(name, n, c, h, w, d, seed, R=None, t=(0.15, 0.02, 0.01), start=0.6, interval=0.05, per_pixel=False, groups=8,
             inv=False):
        rng = np.random.default_rng(seed)
        R = np.eye(3, dtype=np.float32) if R is None else R.astype(np.float32)
        fx = 0.9 * w
        left = _vis_cam(np.eye(3, dtype=np.float32), (0, 0, 0), fx, fx, 0.5 * w, 0.5 * h, n)
        right = _vis_cam(R, t, fx * 1.02, fx * 0.98, 0.49 * w, 0.51 * h, n)
        if per_pixel:
            ds = (start + 0.2 * rng.random((n, 1, h, w))).astype(np.float32)
            di = (interval * (0.5 + rng.random((n, 1, h, w)))).astype(np.float32)
        else:
            ds = np.full((n, 1, 1, 1), start, np.float32) + np.arange(n, dtype=np.float32).reshape(n, 1, 1, 1) * 0.1
            di = np.full((n, 1, 1, 1), interval, np.float32)
        return dict(name=name, ref=syn.features((n, c, h, w), seed), src=syn.features((n, c, h, w), seed + 1), left_cam=left,
                    right_cam=right, depth_num=d, depth_start=ds, depth_interval=di, groups=groups, inv=inv)

    out.append(case("vis_small", 2, 16, 12, 16, 6, 80))
    out.append(case("vis_rot", 1, 32, 20, 28, 8, 81, R=syn.rot_y(4) @ syn.rot_z(7), t=(0.2, -0.05, 0.03)))
    out.append(case("vis_per_pixel", 2, 16, 10, 14, 5, 82, per_pixel=True, R=syn.rot_x(3)))
    out.append(case("vis_inv_depth", 1, 8, 9, 13, 7, 83, inv=True, groups=4))
    out.append(case("vis_behind", 1, 8, 10, 12, 6, 84, R=syn.rot_y(170), t=(0.1, 0, 0.8), groups=2))
    out.append(case("vis_dtu_like", 1, 32, 64, 80, 32, 85, t=(0.12, 0.01, 0.0), start=0.45, interval=0.016))
    return out


def cvp_cases():
    """CVP-MVSNet cost-volume cases: reference + nsrc source maps ``[B,C,H,W]``, pixel intrinsics
    ``[B,3,3]`` / ``[B,nsrc,3,3]``, world-to-camera extrinsics ``[B,4,4]`` / ``[B,nsrc,4,4]``, per-pixel hypotheses
    ``[B,D,H,W]`` (and per-plane ``[B,D]`` for ``homo_warping``)."""
    out = []

    def case(name, B, C, H, W, D, nsrc, seed, rot=None):
        rng = np.random.default_rng(seed)
        K = np.array([[0.9 * W, 0, 0.5 * W], [0, 0.9 * W, 0.5 * H], [0, 0, 1]], np.float32)
        ref_in = np.repeat(K[None], B, 0)
        src_in = np.repeat(np.repeat((K * np.array([[1.02], [0.98], [1.0]], np.float32))[None, None], B, 0), nsrc, 1)
        ref_ex = np.repeat(np.eye(4, dtype=np.float32)[None], B, 0)
        src_ex = np.zeros((B, nsrc, 4, 4), np.float32)
        for v in range(nsrc):
            E = np.eye(4, dtype=np.float32)
            E[:3, :3] = (syn.rot_y(1.5 * (v + 1)) if rot is None else rot(v)).astype(np.float32)
            E[:3, 3] = (0.08 * (v + 1) * (-1) ** v, 0.01 * v, 0.005)
            src_ex[:, v] = E
        base = (0.5 + 0.3 * rng.random((B, 1, H, W))).astype(np.float32)
        hyp = (base + 0.02 * np.arange(D, dtype=np.float32).reshape(1, D, 1, 1)).astype(np.float32)
        planes = np.repeat((0.45 + 0.03 * np.arange(D, dtype=np.float32))[None], B, 0)
        return dict(name=name, ref=syn.features((B, C, H, W), seed), srcs=[syn.features((B, C, H, W), seed + 1 + v) for v in range(nsrc)],
                    ref_in=ref_in, src_in=src_in, ref_ex=ref_ex, src_ex=src_ex, depth_hypos=hyp, depth_planes=planes, nsrc=nsrc)

    out.append(case("cvp_small", 2, 8, 10, 14, 5, 2, 90))
    out.append(case("cvp_odd_c", 1, 5, 9, 11, 4, 3, 91, rot=lambda v: syn.rot_z(6 * (v + 1)) @ syn.rot_x(2)))
    out.append(case("cvp_c16", 1, 16, 24, 32, 8, 4, 92))
    return out
