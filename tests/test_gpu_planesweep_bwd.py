"""GPU parity of the correlation backward (``rmvd_psweep_corr_bwd`` through the autograd node of the
block) against the numpy oracle and the golden gradients frozen from the UNMODIFIED reference's autograd
(tools/make_bwd_golden.py).  Bar: gradients within 1e-4 of max|ref| per tensor (the tolerance BASELINE.json states for the
cost volume); the forward outputs of the differentiable path are bit-identical to the no-grad path."""
import numpy as np
import pytest
import torch

import oracle
from tools import cases
from robustmvd_b200 import synthetic as syn
from golden_util import load_golden

pytestmark = pytest.mark.gpu
BW = cases.backward_cases()
TOL = 1e-4
DEV = "cuda:0"


def _t(a, grad=False):
    if a is None:
        return None
    x = torch.from_numpy(np.asarray(a)).to(DEV)
    return x.requires_grad_(True) if grad else x


def _block_inputs(case, key_grad=True, src_grad=True):
    kw = {k: (_t(v) if isinstance(v, np.ndarray) else v) for k, v in case["kwargs"].items()}
    key = _t(case["feat_key"], key_grad)
    srcs = [_t(f, src_grad) for f in case["feat_sources"]]
    args = dict(feat_key=key, intrinsics_key=_t(case["intrinsics_key"]), feat_sources=srcs,
                source_to_key_transforms=[_t(x) for x in case["transforms"]],
                intrinsics_sources=None if case["intrinsics_sources"] is None else [_t(x) for x in case["intrinsics_sources"]],
                **kw)
    return key, srcs, args


def _run_backward(case, plane_range=None, key_grad=True, src_grad=True, gs=None):
    from robustmvd_b200.blocks.planesweep_corr import PlanesweepCorrelation
    key, srcs, args = _block_inputs(case, key_grad, src_grad)
    blk = PlanesweepCorrelation(normalize=case.get("normalize", "dim"))
    corrs, masks, inv = blk(plane_range=plane_range, **args)
    assert corrs[0].requires_grad and not masks[0].requires_grad
    S = inv.shape[1]
    gs = cases.grad_corrs(case, S) if gs is None else gs
    s0, s1 = (0, S) if plane_range is None else plane_range
    loss = sum((c * _t(np.ascontiguousarray(g[:, s0:s1]))).sum() for c, g in zip(corrs, gs))
    wrt = ([key] if key_grad else []) + (srcs if src_grad else [])
    grads = torch.autograd.grad(loss, wrt)
    torch.cuda.synchronize()
    return [g.cpu().numpy() for g in grads], [c.detach().cpu().numpy() for c in corrs], [m.cpu().numpy() for m in masks], S


def _rel(a, b):
    return float(np.abs(a - b).max()) / max(float(np.abs(b).max()), 1e-30)


@pytest.mark.parametrize("case", BW, ids=[c["name"] for c in BW])
def test_backward_matches_oracle_and_reference_autograd(cuda_lib, case):
    grads, corrs, masks, S = _run_backward(case)
    ogk, ogs = oracle.planesweep_correlation_backward(
        case["feat_key"], case["intrinsics_key"], case["feat_sources"], case["transforms"], cases.grad_corrs(case, S),
        case["intrinsics_sources"], normalize=case.get("normalize", "dim"), **case["kwargs"])
    g = load_golden("psbwd", case["name"])
    assert _rel(grads[0], ogk) <= TOL, f"grad_key vs oracle {_rel(grads[0], ogk)}"
    assert _rel(grads[0], g["grad_key"]) <= TOL, f"grad_key vs reference autograd {_rel(grads[0], g['grad_key'])}"
    for v in range(len(ogs)):
        assert _rel(grads[1 + v], ogs[v]) <= TOL, f"grad_src {v} vs oracle {_rel(grads[1 + v], ogs[v])}"
        assert _rel(grads[1 + v], g[f"grad_src_{v}"]) <= TOL, f"grad_src {v} vs reference autograd"
    # the forward outputs of the differentiable path: masks bit-exact, corr within the forward tolerance
    oc, om, _ = oracle.planesweep_correlation(case["feat_key"], case["intrinsics_key"], case["feat_sources"],
                                              case["transforms"], case["intrinsics_sources"],
                                              normalize=case.get("normalize", "dim"), **case["kwargs"])
    for v in range(len(oc)):
        assert np.array_equal(masks[v], om[v])
        assert _rel(corrs[v], oc[v]) <= TOL


@pytest.mark.parametrize("name", ["bwd_c128_stereo", "bwd_c256_kitti_small", "odd_sizes", "band_srcres_batch2_c32", "bwd_roll40_c32"])
def test_direct_kernel_matches_reference_autograd(cuda_lib, monkeypatch, name):
    """The general direct kernel (C > 256 in production) forced onto shapes the tiled kernel normally takes."""
    monkeypatch.setenv("RMVD_PSWEEP_BWD_KERNEL", "direct")
    case = [c for c in BW if c["name"] == name][0]
    grads, _, _, _ = _run_backward(case)
    g = load_golden("psbwd", name)
    assert _rel(grads[0], g["grad_key"]) <= TOL
    for v in range(len(grads) - 1):
        assert _rel(grads[1 + v], g[f"grad_src_{v}"]) <= TOL


def test_wide_channel_count_takes_the_direct_kernel(cuda_lib):
    """C = 384 > 256: the direct vectorised kernel; checked against the oracle."""
    T = syn.make_transform
    case = cases._ps("bwd_c384", 6, 12, 384, [T(syn.rot_y(2), (0.3, 0, 0.02))], S=8, seed=91)
    grads, _, _, S = _run_backward(case)
    ogk, ogs = oracle.planesweep_correlation_backward(
        case["feat_key"], case["intrinsics_key"], case["feat_sources"], case["transforms"], cases.grad_corrs(case, S),
        **case["kwargs"])
    assert _rel(grads[0], ogk) <= TOL and _rel(grads[1], ogs[0]) <= TOL


def test_only_requested_gradients_are_computed(cuda_lib):
    case = [c for c in BW if c["name"] == "bwd_c128_stereo"][0]
    full, _, _, _ = _run_backward(case)
    only_key, _, _, _ = _run_backward(case, src_grad=False)
    only_src, _, _, _ = _run_backward(case, key_grad=False)
    assert len(only_key) == 1 and _rel(only_key[0], full[0]) <= 1e-6
    assert len(only_src) == 2
    for v in range(2):
        assert _rel(only_src[v], full[1 + v]) <= 1e-5   # float atomics: summation order differs between launches


def test_depth_slab_gradients_add_up(cuda_lib):
    """plane_range slabs: the gradients of two depth slabs sum to the gradients of the full sweep."""
    case = [c for c in BW if c["name"] == "band_stereo_c32"][0]
    full, _, _, S = _run_backward(case)
    a, _, _, _ = _run_backward(case, plane_range=(0, 23))
    b, _, _, _ = _run_backward(case, plane_range=(23, S))
    for f, x, y in zip(full, a, b):
        assert _rel(x + y, f) <= 1e-5


def test_more_views_than_one_launch_takes(cuda_lib):
    T = syn.make_transform
    views = [T(syn.rot_y(0.5 * k - 4), (0.05 * k - 0.4, 0.01 * k, 0.0)) for k in range(18)]
    case = cases._ps("bwd_18_views", 8, 10, 8, views, S=6, seed=90)
    grads, _, _, S = _run_backward(case)
    ogk, ogs = oracle.planesweep_correlation_backward(
        case["feat_key"], case["intrinsics_key"], case["feat_sources"], case["transforms"], cases.grad_corrs(case, S),
        **case["kwargs"])
    assert _rel(grads[0], ogk) <= TOL
    for v in range(18):
        assert _rel(grads[1 + v], ogs[v]) <= TOL


def test_adjoint_identity_at_full_kitti_shape(cuda_lib):
    """BASELINE configs[1] at full size (48x160 features, C=256, S=256, 2 source views): corr is linear in the key map and
    in each source map, so <g, corr_v> = <d src_v, src_v> and sum_v <g_v, corr_v> = <d key, key>; and the backward is linear
    in g.  Size-independent properties (the oracle would need minutes at this size)."""
    from robustmvd_b200.blocks.planesweep_corr import PlanesweepCorrelation
    c = syn.CFG2
    inp = syn.planesweep_inputs(c["h"], c["w"], c["C"], 2, n=1, seed=5)
    key = _t(inp["feat_key"], True)
    srcs = [_t(f, True) for f in inp["feat_sources"]]
    blk = PlanesweepCorrelation()
    kw = dict(intrinsics_key=_t(inp["intrinsics_key"]), source_to_key_transforms=[_t(x) for x in inp["transforms"]],
              num_sampling_points=c["S"], min_depth=c["min_depth"], max_depth=c["max_depth"])
    corrs, masks, _ = blk(feat_key=key, feat_sources=srcs, **kw)
    gen = torch.Generator(device=DEV).manual_seed(11)
    g1 = [torch.randn(x.shape, device=DEV, generator=gen) for x in corrs]
    g2 = [torch.randn(x.shape, device=DEV, generator=gen) for x in corrs]
    d1 = torch.autograd.grad(corrs, [key] + srcs, g1, retain_graph=True)
    d2 = torch.autograd.grad(corrs, [key] + srcs, g2, retain_graph=True)
    d3 = torch.autograd.grad(corrs, [key] + srcs, [2.5 * a - b for a, b in zip(g1, g2)])
    f64 = lambda x: x.detach().double()  # noqa: E731
    lhs = [float((f64(g) * f64(cv)).sum()) for g, cv in zip(g1, corrs)]
    scale = sum(float((f64(g).abs() * f64(cv).abs()).sum()) for g, cv in zip(g1, corrs))
    assert abs(sum(lhs) - float((f64(d1[0]) * f64(key)).sum())) <= 1e-5 * scale
    for v in range(2):
        assert abs(lhs[v] - float((f64(d1[1 + v]) * f64(srcs[v])).sum())) <= 1e-5 * scale
    for a, b, c3 in zip(d1, d2, d3):
        ref = 2.5 * a - b
        assert float((c3 - ref).abs().max()) <= 1e-4 * float(ref.abs().max())
    # a masked-out sample passes no gradient: zero upstream gradient wherever the mask is 1 -> all gradients are zero
    gz = [torch.randn(x.shape, device=DEV, generator=gen) * (1 - m) for x, m in zip(corrs, masks)]
    corrs, masks, _ = blk(feat_key=key, feat_sources=srcs, **kw)
    dz = torch.autograd.grad(corrs, [key] + srcs, gz)
    assert all(float(x.abs().max()) == 0.0 for x in dz)


def test_backward_rejects_unsupported_mode(cuda_lib):
    import ctypes
    one = ctypes.c_void_p(256)
    arr = (ctypes.c_void_p * 1)(256)
    ints = (ctypes.c_int * 1)(4)
    st = (ctypes.c_int64 * 4)(0, 1, 0, 0)
    rc = cuda_lib.rmvd_psweep_corr_bwd(one, arr, one, arr, arr, one, st, 1, 8, 4, 4, 1, ints, ints, 4, 0, 4, 2, one, 4, 0,
                                       one, arr, one, 1 << 20, None)
    assert rc == -3   # RMVD_E_MODE: L2-normalised inputs are differentiated around a NORM_NONE call


@pytest.mark.parametrize("case", cases.warp_backward_cases(), ids=lambda c: c["name"])
def test_warp_only_backward_matches_reference_autograd(cuda_lib, case):
    """Training direction of the warp-only block (WarpOnlyCorr, planesweep_corr.py:107-140): autograd through
    PlanesweepCorrelation(warp_only=True) -- rmvd_psweep_warp_fwd / rmvd_psweep_warp_bwd, the L2 normalisations as torch ops around
    them -- against gradients frozen from the UNMODIFIED reference block under CPU autograd (tools/make_warp_bwd_golden.py).
    Tolerance 1e-4 of max|grad| per tensor; the forward values equal the forward-only call."""
    import os
    from robustmvd_b200.blocks.planesweep_corr import PlanesweepCorrelation
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", f"warpbwd_{case['name']}.npz"))
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
    blk = PlanesweepCorrelation(warp_only=True, normalize=case["normalize"])
    srcs = [t(f).requires_grad_(True) for f in case["feat_sources"]]
    kw = dict(feat_key=t(case["feat_key"]), intrinsics_key=t(case["intrinsics_key"]),
              source_to_key_transforms=[t(x) for x in case["transforms"]], **case["kwargs"])
    warped, masks, _ = blk(feat_sources=srcs, **kw)
    assert all(wv.requires_grad for wv in warped) and not any(m.requires_grad for m in masks)
    ups = [t(u) for u in cases.warp_upstream_grads(case, tuple(warped[0].shape))]
    grads = torch.autograd.grad(sum((wv * u).sum() for wv, u in zip(warped, ups)), srcs)
    for v, got in enumerate(grads):
        ref = g[f"grad_src_{v}"]
        assert np.abs(got.cpu().numpy() - ref).max() <= 1e-4 * np.abs(ref).max(), v
    with torch.no_grad():
        w0, m0, _ = PlanesweepCorrelation(warp_only=True, normalize=case["normalize"])(feat_sources=[s.detach() for s in srcs], **kw)
    for a, b, ma, mb in zip(warped, w0, masks, m0):
        assert torch.equal(ma, mb) and float((a.detach() - b).abs().max()) <= 1e-5 * float(b.abs().max())
