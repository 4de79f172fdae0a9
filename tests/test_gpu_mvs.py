"""GPU parity: MVSNet variance volume, homo_warp, softmax regression + confidence, slab softargmax, soft_argmin and
entropy (through the C ABI) against the numpy oracle and the golden vectors of the unmodified reference.
Bars: sampling indices bit-exact, cost volumes within 1e-4 of max|ref|, depth within 1e-3 relative."""
import numpy as np
import pytest
import torch

import oracle
from tools import cases
from robustmvd_b200 import synthetic as syn
from golden_util import (load_golden, check_mvs_variance_against_golden, check_warp_against_golden,
                         check_regress_against_golden)

pytestmark = pytest.mark.gpu
MVS = cases.mvs_cases(include_large=True)
cu = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
npy = lambda t: t.cpu().numpy()  # noqa: E731


@pytest.mark.parametrize("case", MVS, ids=[c["name"] for c in MVS])
def test_variance_and_warp(cuda_lib, case):
    from robustmvd_b200.blocks import utils as bu
    g = load_golden("mvs", case["name"])
    var = npy(bu.mvs_variance_volume(cu(case["feats"]), cu(case["proj"]), cu(case["depth"])))
    ovar = oracle.mvs_variance(case["feats"], case["proj"], case["depth"])
    assert np.abs(var - ovar).max() <= 1e-4 * np.abs(ovar).max()
    check_mvs_variance_against_golden(g, var)
    w1, ix0, iy0 = bu.homo_warp(cu(case["feats"][:, 1]), cu(case["proj"][:, 1]), cu(case["proj"][:, 0]),
                                cu(case["depth"]), return_indices=True)
    ow, odbg = oracle.homo_warp(case["feats"][:, 1], case["proj"][:, 1], case["proj"][:, 0], case["depth"],
                                return_debug=True)
    assert np.array_equal(npy(ix0), odbg["ix0"]) and np.array_equal(npy(iy0), odbg["iy0"])
    assert np.abs(npy(w1) - ow).max() <= 1e-4 * np.abs(ow).max()
    check_warp_against_golden(g, npy(w1))
    # depth slabs of the variance volume concatenate to the full volume, bit for bit
    D = case["depth"].shape[1]
    cut = D // 3 + 1
    parts = [npy(bu.mvs_variance_volume(cu(case["feats"]), cu(case["proj"]), cu(case["depth"]), plane_range=r))
             for r in ((0, cut), (cut, D))]
    assert np.array_equal(np.concatenate(parts, 2), var)


@pytest.mark.parametrize("case", MVS, ids=[c["name"] for c in MVS])
def test_regression(cuda_lib, case):
    from robustmvd_b200.blocks import utils as bu
    g = load_golden("mvs", case["name"])
    r = bu.mvs_regress(cu(case["cost"]), cu(case["depth"]), return_prob=True)
    depth, conf, didx, eidx = npy(r["depth"]), npy(r["confidence"]), npy(r["depth_index"]), npy(r["expected_index"])
    check_regress_against_golden(g, depth, conf, didx, eidx)
    od, oc, oi, oe = oracle.mvs_regress(case["cost"], case["depth"])
    assert (np.abs(depth - od) <= 1e-3 * np.abs(od)).all()
    p = npy(r["prob"])
    assert np.abs(p.sum(1) - 1).max() < 1e-5
    # depth_regression() on the probabilities reproduces the fused result
    d2 = npy(bu.depth_regression(r["prob"], cu(case["depth"])))
    assert (np.abs(d2 - od) <= 1e-3 * np.abs(od)).all()
    prob, sa, pm = bu.soft_argmin(cu(case["cost"]), 1, keepdim=True, window=2)
    assert np.abs(npy(sa) - g["soft_argmin"]).max() < 1e-3
    near = np.abs(np.abs(np.arange(p.shape[1], dtype=np.float32).reshape(1, -1, 1, 1) - g["soft_argmin"]) - 2) < 1e-3
    assert np.abs(npy(pm) - g["prob_map"])[~near.any(1, keepdims=True)].max() < 1e-4
    ent = npy(bu.entropy(prob, 1, keepdim=True))
    assert np.abs(ent - g["entropy"]).max() < 1e-4


@pytest.mark.parametrize("G", [2, 3, 8])
def test_slab_softargmax_equals_full(cuda_lib, G):
    """Depth-slab partition of the regression: partial -> combine -> confidence parts summed."""
    from robustmvd_b200 import _lib
    lib = cuda_lib
    case = MVS[-1]
    cost, depth = case["cost"], case["depth"]
    B, D, h, w = cost.shape
    od, oc, oi, oe = oracle.mvs_regress(cost, depth)
    bounds = np.linspace(0, D, G + 1).astype(int)
    parts = torch.empty((G, 4, B, h, w), dtype=torch.float32, device="cuda")
    slabs = []
    for gi in range(G):
        lo, hi = int(bounds[gi]), int(bounds[gi + 1])
        cs, ds = cu(cost[:, lo:hi]), cu(depth[:, lo:hi])
        slabs.append((cs, lo, hi))
        _lib.check(lib.rmvd_softargmax_partial(_lib.ptr(cs), _lib.ptr(ds), B, hi - lo, h, w, lo, _lib.ptr(parts[gi]),
                                               _lib.stream_ptr()), "partial")
    merged = torch.empty((4, B, h, w), dtype=torch.float32, device="cuda")
    _lib.check(lib.rmvd_softargmax_combine(_lib.ptr(parts), G, B, h, w, _lib.ptr(merged), _lib.stream_ptr()), "combine")
    conf = torch.zeros((B, h, w), dtype=torch.float32, device="cuda")
    for cs, lo, hi in slabs:
        part = torch.empty_like(conf)
        _lib.check(lib.rmvd_softargmax_confidence(_lib.ptr(cs), _lib.ptr(merged), B, hi - lo, h, w, D, lo,
                                                  _lib.ptr(part), _lib.stream_ptr()), "confidence")
        conf += part
    m = npy(merged)
    assert (np.abs(m[2] - od) <= 1e-3 * np.abs(od)).all()
    near = np.abs(oe - np.round(oe)) < 1e-3
    same = m[3].astype(np.int64).clip(0, D - 1) == oi
    assert (same | near).all()
    assert np.abs(npy(conf) - oc)[same].max() < 1e-4
    # oracle partials agree too
    op = oracle.softargmax_partials(cost[:, bounds[0]:bounds[1]], depth[:, bounds[0]:bounds[1]], 0)
    p0 = npy(parts[0])
    assert np.array_equal(p0[0], op[0]) and np.abs(p0[1] - op[1]).max() <= 1e-4 * np.abs(op[1]).max()


def test_cfg3_full_size_properties(cuda_lib):
    """BASELINE.json config 3 at full size (DTU 224x304, C=32, 5 views, D=192): size-independent properties.
    (1) identical source views with an identity homography give zero variance; (2) the regression of a one-hot
    cost recovers the plane's depth; (3) a strided subsample of the variance matches the oracle computed on the
    same voxels' rows."""
    from robustmvd_b200.blocks import utils as bu
    cfg = syn.CFG3
    inp = syn.mvs_inputs(cfg["h"], cfg["w"], cfg["C"], cfg["V"], cfg["D"])
    feats = inp["feats"].copy()
    feats[:, 1:] = feats[:, :1]
    proj = inp["proj"].copy()
    eye = np.eye(4, dtype=np.float32)
    proj[:] = eye
    var = bu.mvs_variance_volume(cu(feats), cu(proj), cu(inp["depth"]))
    # identity homography: ix = x * W/(W-1) - 0.5 is not the pixel centre, so compare against homo_warp itself
    w1 = bu.homo_warp(cu(feats[:, 1]), cu(proj[:, 1]), cu(proj[:, 0]), cu(inp["depth"]))
    ref = cu(feats[:, 0]).unsqueeze(2)
    V = cfg["V"]
    s = ref + (V - 1) * w1
    q = ref ** 2 + (V - 1) * w1 ** 2
    expect = q / V - (s / V) ** 2
    assert float((var - expect).abs().max()) <= 1e-4 * float(expect.abs().max())
    del var, w1, s, q, expect
    # full-size variance on the real geometry: rows 0, 97 and 223 against the oracle restricted to those rows is not
    # possible (warps read other rows), so check a 24-plane slab of a cropped problem instead
    var_full = bu.mvs_variance_volume(cu(inp["feats"]), cu(inp["proj"]), cu(inp["depth"]), plane_range=(100, 104))
    ov = oracle.mvs_variance(inp["feats"], inp["proj"], inp["depth"][:, 100:104])
    assert np.abs(npy(var_full) - ov).max() <= 1e-4 * np.abs(ov).max()
    # one-hot cost
    D = cfg["D"]
    cost = np.full((1, D, cfg["h"], cfg["w"]), -80.0, np.float32)
    k = np.random.default_rng(3).integers(2, D - 3, (cfg["h"], cfg["w"]))
    np.put_along_axis(cost, k[None, None], 80.0, 1)
    r = bu.mvs_regress(cu(cost), cu(inp["depth"]))
    assert np.array_equal(npy(r["depth_index"])[0], k)
    assert np.allclose(npy(r["depth"])[0], inp["depth"][0][k], rtol=1e-6)
    assert np.allclose(npy(r["confidence"]), 1.0, atol=1e-6)


@pytest.mark.parametrize("w", [76, 78], ids=["w%4==0 (16-byte staging)", "w%4!=0 (4-byte staging)"])
def test_variance_rotated_views(cuda_lib, w):
    """Strong in-plane rotation between the views: the bounding rectangle of a tile's samples exceeds the staged
    capacity for some tiles (direct gather path) and fits for others; both must agree with the oracle."""
    from robustmvd_b200.blocks import utils as bu
    inp = syn.mvs_inputs(56, w, 32, 4, 12, seed=77, fx=90.0)
    K = np.array([[90.0, 0, w / 2.0, 0], [0, 90.0, 28.0, 0], [0, 0, 1, 0], [0, 0, 0, 1]], np.float64)
    for v, roll in ((1, 35.0), (2, -80.0), (3, 8.0)):
        pose = np.eye(4)
        pose[:3, :3] = syn.rot_z(roll) @ syn.rot_y(2.0 * v)
        pose[:3, 3] = (0.05 * v, -0.03, 0.01)
        P = pose.astype(np.float32)
        P[:3, :4] = (K[:3, :3].astype(np.float32) @ P[:3, :4]).astype(np.float32)
        inp["proj"][:, v] = P
    var = npy(bu.mvs_variance_volume(cu(inp["feats"]), cu(inp["proj"]), cu(inp["depth"])))
    ov = oracle.mvs_variance(inp["feats"], inp["proj"], inp["depth"])
    assert np.abs(var - ov).max() <= 1e-4 * np.abs(ov).max()


CL_CASES = {
    # name: (h, w, C, views, planes, batch, fx, seed)
    "c32_odd_sizes": (21, 37, 32, 3, 9, 1, 30.0, 81),          # partially filled 16 x 4 tiles on both edges, ragged plane group
    "c64_two_channel_groups": (24, 40, 64, 4, 10, 1, 32.0, 82),
    "c32_batch2_single_source": (18, 34, 32, 2, 7, 2, 28.0, 83),
    "c32_many_views": (16, 32, 32, 9, 5, 1, 25.0, 84),
    "c32_one_plane": (12, 20, 32, 3, 1, 1, 20.0, 85),
}


@pytest.mark.parametrize("name", list(CL_CASES))
def test_variance_channels_last_kernel(cuda_lib, name):
    """Shapes that take the channels-last staged kernel (csrc/mvs_var.cu, C % 32 == 0) against the oracle: ragged tiles, several
    channel groups, batches, one source view, more views than the pipeline has buffers, a single plane."""
    from robustmvd_b200.blocks import utils as bu
    h, w, C, V, D, n, fx, seed = CL_CASES[name]
    inp = syn.mvs_inputs(h, w, C, V, D, n=n, seed=seed, fx=fx)
    var = npy(bu.mvs_variance_volume(cu(inp["feats"]), cu(inp["proj"]), cu(inp["depth"])))
    ov = oracle.mvs_variance(inp["feats"], inp["proj"], inp["depth"])
    assert var.shape == ov.shape
    assert np.abs(var - ov).max() <= 1e-4 * np.abs(ov).max()
    if D > 2:
        part = npy(bu.mvs_variance_volume(cu(inp["feats"]), cu(inp["proj"]), cu(inp["depth"]), plane_range=(1, D - 1)))
        assert np.array_equal(part, var[:, :, 1:D - 1])


def test_variance_views_outside_and_zoomed(cuda_lib):
    """A source view that sees nothing of the key frustum (every sample outside: zeros, mvsnet.py:128 adds them), one zoomed
    2x (neighbouring key pixels two texels apart) and one zoomed out (neighbouring pixels share texels); and a non-finite
    depth plane (nan coordinates sample nothing)."""
    from robustmvd_b200.blocks import utils as bu
    h, w = 20, 36
    inp = syn.mvs_inputs(h, w, 32, 4, 8, seed=91, fx=30.0)
    K = np.array([[30.0, 0, w / 2.0, 0], [0, 30.0, h / 2.0, 0], [0, 0, 1, 0], [0, 0, 0, 1]], np.float64)
    for v, (scale, shift) in ((1, (1.0, 500.0)), (2, (2.0, 0.02)), (3, (0.4, -0.03))):
        Kv = K.copy()
        Kv[0, 0] *= scale
        Kv[1, 1] *= scale
        pose = np.eye(4)
        pose[:3, 3] = (shift, 0.01, 0.0)
        P = pose.astype(np.float32)
        P[:3, :4] = (Kv[:3, :3].astype(np.float32) @ P[:3, :4]).astype(np.float32)
        inp["proj"][:, v] = P
    var = npy(bu.mvs_variance_volume(cu(inp["feats"]), cu(inp["proj"]), cu(inp["depth"])))
    ov = oracle.mvs_variance(inp["feats"], inp["proj"], inp["depth"])
    assert np.abs(var - ov).max() <= 1e-4 * np.abs(ov).max()
    depth = inp["depth"].copy()
    depth[:, 3] = np.nan
    var = npy(bu.mvs_variance_volume(cu(inp["feats"]), cu(inp["proj"]), cu(depth)))
    ov = oracle.mvs_variance(inp["feats"], inp["proj"], depth)
    assert np.abs(var - ov).max() <= 1e-4 * np.abs(ov).max()


def test_variance_kernels_agree_and_workspace_is_checked(cuda_lib):
    """rmvd_mvs_variance_fwd (NCHW-staged kernel) and rmvd_mvs_variance_fwd_ws (channels-last staged kernel) on the same
    inputs; a missing or short workspace is refused."""
    from robustmvd_b200 import _lib
    lib = cuda_lib
    inp = syn.mvs_inputs(28, 44, 32, 4, 12, seed=95, fx=40.0)
    feats, proj, depth = cu(inp["feats"]), cu(inp["proj"]), cu(inp["depth"])
    B, V, C, h, w = feats.shape
    D = depth.shape[1]
    a = torch.empty((B, C, D, h, w), dtype=torch.float32, device="cuda")
    b = torch.empty_like(a)
    need = lib.rmvd_mvs_variance_workspace_bytes(B, V, C, h, w)
    assert need == 4 * B * (V - 1) * C * h * w
    assert lib.rmvd_mvs_variance_workspace_bytes(B, V, 24, h, w) == 0 and lib.rmvd_mvs_variance_workspace_bytes(B, 1, C, h, w) == 0
    ws = torch.empty(need, dtype=torch.uint8, device="cuda")
    args = (_lib.ptr(feats), _lib.ptr(proj), _lib.ptr(depth), B, V, C, h, w, D, 0, D)
    _lib.check(lib.rmvd_mvs_variance_fwd(*args, _lib.ptr(a), D, 0, _lib.stream_ptr()), "rmvd_mvs_variance_fwd")
    _lib.check(lib.rmvd_mvs_variance_fwd_ws(*args, _lib.ptr(b), D, 0, _lib.ptr(ws), need, _lib.stream_ptr()), "rmvd_mvs_variance_fwd_ws")
    assert float((a - b).abs().max()) <= 2e-5 * float(a.abs().max())
    assert lib.rmvd_mvs_variance_fwd_ws(*args, _lib.ptr(b), D, 0, None, need, _lib.stream_ptr()) == -1          # RMVD_E_NULL
    assert lib.rmvd_mvs_variance_fwd_ws(*args, _lib.ptr(b), D, 0, _lib.ptr(ws), need - 1, _lib.stream_ptr()) == -4  # RMVD_E_WORKSPACE


def test_cfg3_full_size_against_reference_golden(cuda_lib):
    """BASELINE.json configs[2] at full size (DTU 224x304 features, C=32, 5 views, 192 planes) against the golden frozen from the
    unmodified reference (tools/make_fullsize_golden.py): a stride-997 subsample of the 418 M-element variance volume within 1e-4 of
    max|ref|, regressed depth within 1e-3, the integer depth index exact away from ties."""
    from robustmvd_b200.blocks import utils as bu
    case = cases.fullsize_mvs_case()
    g = load_golden("mvs", case["name"])
    var = bu.mvs_variance_volume(cu(case["feats"]), cu(case["proj"]), cu(case["depth"]))
    sub = int(g["sub"])
    got = npy(var.reshape(-1)[::sub])
    assert np.abs(got - g["var_sub"]).max() <= 1e-4 * float(g["var_absmax"])
    del var
    r = bu.mvs_regress(cu(case["cost"]), cu(case["depth"]))
    check_regress_against_golden(g, npy(r["depth"]), npy(r["confidence"]), npy(r["depth_index"]), npy(r["expected_index"]))


def test_depth_regression_and_entropy_follow_the_reference_for_any_input(cuda_lib):
    """blocks/utils.py:271-274 and :67-68 apply no softmax: an un-normalised volume, per-pixel depth hypotheses
    (cvp_mvsnet.py:168) and values outside the clamp range of entropy() must come out as torch computes them."""
    from robustmvd_b200.blocks import utils as bu
    from robustmvd_b200._lib import RmvdError
    rng = np.random.default_rng(5)
    p = (rng.random((2, 9, 7, 13), dtype=np.float32) * 1.7).astype(np.float32)          # not normalised, some values > 1
    p[0, 3] = 0.0
    dv = rng.random((2, 9), dtype=np.float32) + 0.5
    dpix = rng.random((2, 9, 7, 13), dtype=np.float32) + 0.5
    tp = torch.from_numpy(p)
    ref = torch.sum(tp * torch.from_numpy(dv).view(2, 9, 1, 1), 1).numpy()
    assert np.allclose(npy(bu.depth_regression(cu(p), cu(dv))), ref, rtol=1e-6, atol=1e-6)
    ref4 = torch.sum(tp * torch.from_numpy(dpix), 1).numpy()
    assert np.allclose(npy(bu.depth_regression(cu(p), cu(dpix))), ref4, rtol=1e-6, atol=1e-6)
    ent = torch.sum(-tp * tp.clamp(1e-9, 1.0).log(), dim=1, keepdim=True).numpy()
    assert np.allclose(npy(bu.entropy(cu(p), 1, keepdim=True)), ent, rtol=1e-5, atol=1e-5)
    with pytest.raises(RmvdError):                                                       # forward only: no silent graph-less result
        bu.depth_regression(cu(p).requires_grad_(True), cu(dv))
    with torch.no_grad():
        bu.depth_regression(cu(p).requires_grad_(True), cu(dv))


@pytest.mark.parametrize("case", cases.mvs_cases(include_large=False), ids=lambda c: c["name"])
def test_mvs_backward_matches_reference_autograd(cuda_lib, case):
    """Training direction of the MVSNet path: autograd through blocks.utils.mvs_variance_volume / mvs_regress (rmvd_mvs_variance_bwd,
    rmvd_mvs_regress_bwd) against gradients frozen from the UNMODIFIED reference functions under CPU autograd
    (tools/make_mvs_bwd_golden.py) and against the numpy oracle.  Tolerance 1e-4 of max|grad| (float atomics sum in any order)."""
    import os
    from robustmvd_b200.blocks import utils as bu
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", f"mvsbwd_{case['name']}.npz"))
    g_var, g_depth = cases.mvs_upstream_grads(case)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
    feats = t(case["feats"]).requires_grad_(True)
    var = bu.mvs_variance_volume(feats, t(case["proj"]), t(case["depth"]))
    (gf,) = torch.autograd.grad((var * t(g_var)).sum(), [feats])
    cost = t(case["cost"]).requires_grad_(True)
    reg = bu.mvs_regress(cost, t(case["depth"]))
    assert not reg["confidence"].requires_grad and reg["depth"].requires_grad
    (gc,) = torch.autograd.grad((reg["depth"] * t(g_depth)).sum(), [cost])
    for got, key, ora in ((gf, "grad_feats", oracle.mvs_variance_backward(case["feats"], case["proj"], case["depth"], g_var)),
                          (gc, "grad_cost", oracle.mvs_regress_backward(case["cost"], case["depth"], g_depth))):
        got = got.cpu().numpy()
        assert np.abs(got - g[key]).max() <= 1e-4 * np.abs(g[key]).max(), key
        assert np.abs(got - ora).max() <= 1e-4 * np.abs(ora).max(), key
    # forward values of the differentiable wrappers are the forward-only ones
    with torch.no_grad():
        assert torch.equal(var, bu.mvs_variance_volume(feats.detach(), t(case["proj"]), t(case["depth"])))


def test_mvs_backward_at_a_larger_shape_against_torch_autograd(cuda_lib):
    """DTU-quarter shape (56 x 76, C = 32, 5 views, 48 planes): the CUDA backward against torch's own autograd through the reference's
    operator sequence run on the same device (oracle/torch_cpu.py with device='cuda': grid_sample + the variance expression)."""
    from oracle import torch_cpu
    from robustmvd_b200.blocks import utils as bu
    case = cases.mvs_cases(include_large=True)[-1]
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
    proj, depth = t(case["proj"]), t(case["depth"])
    gen = torch.Generator(device="cuda").manual_seed(7)
    feats = t(case["feats"]).requires_grad_(True)
    var = bu.mvs_variance_volume(feats, proj, depth)
    g_var = torch.randn(var.shape, device="cuda", generator=gen)
    (gf,) = torch.autograd.grad((var * g_var).sum(), [feats])
    feats2 = t(case["feats"]).requires_grad_(True)
    var2 = torch_cpu.mvs_variance(feats2, proj, depth, device="cuda", return_torch=True)
    (gf2,) = torch.autograd.grad((var2 * g_var).sum(), [feats2])
    assert float((gf - gf2).abs().max()) <= 2e-4 * float(gf2.abs().max())


def test_mvsnet_trains_through_the_cuda_path(cuda_lib):
    """``mvsnet_train`` under train=True: one forward + backward through FeatureNet -> variance volume (CUDA) -> CostRegNet ->
    regression (CUDA) produces finite, non-zero gradients for the first and the last layer."""
    import robustmvd_b200 as rb
    torch.manual_seed(cases.MODEL_SEED)
    model = rb.create_model("mvsnet_train", pretrained=False, num_gpus=1, train=True)
    cases.prepare_model_for_case(model, "mvsnet_train")
    sample = model.input_adapter(**syn.model_sample(64, 96, 2, seed=13, batch=1, depth_range=(2.0, 40.0)))
    pred, _ = model(**sample)
    pred["depth"].mean().backward()
    grads = {k: p.grad for k, p in model.named_parameters() if p.grad is not None}
    first = next(k for k in grads if k.startswith("feature"))
    assert len(grads) > 10 and all(bool(torch.isfinite(v).all()) for v in grads.values())
    assert float(grads[first].abs().max()) > 0 and float(grads["cost_regularization.prob.weight"].abs().max()) > 0
