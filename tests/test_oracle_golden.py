"""CPU: the numpy oracle against the golden vectors frozen from the UNMODIFIED reference (tools/make_golden.py)."""
import os

import numpy as np
import pytest

import oracle
from tools import cases
from golden_util import (load_golden, check_planesweep_against_golden, check_mvs_variance_against_golden,
                         check_warp_against_golden, check_regress_against_golden)

PS = cases.planesweep_cases(include_large=True)
MVS = cases.mvs_cases(include_large=True)


@pytest.mark.parametrize("case", PS, ids=[c["name"] for c in PS])
def test_planesweep_oracle_matches_golden(case):
    g = load_golden("ps", case["name"])
    corrs, masks, inv, dbg = oracle.planesweep_correlation(
        case["feat_key"], case["intrinsics_key"], case["feat_sources"], case["transforms"], case["intrinsics_sources"],
        warp_only=case.get("warp_only", False), normalize=case.get("normalize", "dim"), return_debug=True,
        **case["kwargs"])
    check_planesweep_against_golden(g, corrs, masks, inv, [d["ix0"] for d in dbg], [d["iy0"] for d in dbg],
                                    [d["geom_mask"] for d in dbg], tol=1e-5)


@pytest.mark.parametrize("case", MVS, ids=[c["name"] for c in MVS])
def test_mvs_oracle_matches_golden(case):
    g = load_golden("mvs", case["name"])
    var = oracle.mvs_variance(case["feats"], case["proj"], case["depth"])
    check_mvs_variance_against_golden(g, var, tol=1e-5)
    w1 = oracle.homo_warp(case["feats"][:, 1], case["proj"][:, 1], case["proj"][:, 0], case["depth"])
    check_warp_against_golden(g, w1, tol=1e-5)
    d, c, i, e = oracle.mvs_regress(case["cost"], case["depth"])
    check_regress_against_golden(g, d, c, i, e, rtol=1e-5)
    p, sa, pm = oracle.soft_argmin(case["cost"], 1, keepdim=True, window=2)
    assert np.abs(sa - g["soft_argmin"]).max() < 1e-3
    assert np.abs(pm - g["prob_map"]).max() < 1e-4
    assert np.abs(oracle.entropy(p, 1, keepdim=True) - g["entropy"]).max() < 1e-4


def test_slab_partials_combine_to_full_regression():
    c = MVS[0]
    D = c["cost"].shape[1]
    d, conf, idx, eidx = oracle.mvs_regress(c["cost"], c["depth"])
    parts = []
    for lo, hi in ((0, 3), (3, D)):
        parts.append(oracle.softargmax_partials(c["cost"][:, lo:hi], c["depth"][:, lo:hi], lo))
    _, _, dep, ei = oracle.softargmax_combine(parts)
    assert np.allclose(dep, d, rtol=1e-5) and np.allclose(ei, eidx, rtol=1e-5, atol=1e-5)


def test_torch_cpu_baseline_port_matches_oracle():
    """The multi-threaded CPU baseline that bench.py times (oracle/torch_cpu.py) computes the same thing."""
    from oracle import torch_cpu
    case = [c for c in PS if c["name"] == "kitti_like_c32"][0]
    g = load_golden("ps", case["name"])
    corrs, masks, inv = torch_cpu.planesweep_correlation(
        case["feat_key"], case["intrinsics_key"], case["feat_sources"], case["transforms"], **case["kwargs"])
    check_planesweep_against_golden(g, corrs, masks, inv, tol=1e-5)
    m = MVS[0]
    var = torch_cpu.mvs_variance(m["feats"], m["proj"], m["depth"])
    check_mvs_variance_against_golden(load_golden("mvs", m["name"]), var, tol=1e-5)


def test_fusion_oracle_against_golden():
    """oracle.view_fusion against the outputs of the unmodified reference LearnedFusion (tools/make_fusion_golden.py)."""
    import glob
    import os
    from golden_util import GOLDEN
    files = sorted(glob.glob(os.path.join(GOLDEN, "fusion_*.npz")))
    assert len(files) >= 4
    for f in files:
        g = np.load(f)
        V = int(g["V"])
        w, fc, fm = oracle.view_fusion([g[f"logit_{v}"] for v in range(V)], [g[f"corr_{v}"] for v in range(V)],
                                       [g[f"mask_{v}"].astype(np.float32) for v in range(V)])
        assert np.array_equal(fm, g["fused_mask"])
        assert np.abs(fc - g["fused_corr"]).max() <= 1e-6 * max(float(np.abs(g["fused_corr"]).max()), 1e-30)
        for v in range(V if V > 1 else 0):
            assert np.abs(w[v] - g[f"weight_{v}"]).max() <= 1e-6


BW = cases.backward_cases()


@pytest.mark.parametrize("case", BW, ids=[c["name"] for c in BW])
def test_backward_oracle_matches_reference_autograd(case):
    """oracle.planesweep_correlation_backward against the gradients of the unmodified reference block
    (tools/make_bwd_golden.py: torch autograd through TorchCorr.forward on CPU)."""
    g = load_golden("psbwd", case["name"])
    S = int(np.asarray(oracle.planesweep_correlation(case["feat_key"], case["intrinsics_key"], case["feat_sources"][:1],
                                                     case["transforms"][:1], **case["kwargs"])[2]).shape[1])
    gk, gs = oracle.planesweep_correlation_backward(
        case["feat_key"], case["intrinsics_key"], case["feat_sources"], case["transforms"], cases.grad_corrs(case, S),
        case["intrinsics_sources"], normalize=case.get("normalize", "dim"), **case["kwargs"])
    assert np.abs(gk - g["grad_key"]).max() <= 1e-5 * np.abs(g["grad_key"]).max()
    for v, x in enumerate(gs):
        assert np.abs(x - g[f"grad_src_{v}"]).max() <= 1e-5 * max(np.abs(g[f"grad_src_{v}"]).max(), 1e-30)


VIS = cases.vis_cases()


@pytest.mark.parametrize("case", VIS, ids=[c["name"] for c in VIS])
def test_vis_cost_volume_oracle_matches_reference(case):
    """oracle.groupwise_cost_volume / homography_warp_volume against the unmodified reference functions
    (tools/make_vis_golden.py: get_homographies + homography_warping + groupwise_correlation on CPU)."""
    g = load_golden("vis", case["name"])
    oc = oracle.groupwise_cost_volume(case["ref"], case["src"], g["Hs"], case["groups"])
    ow = oracle.homography_warp_volume(case["src"], g["Hs"])
    assert np.abs(oc.reshape(-1)[::int(g["cost_stride"])] - g["cost_sub"]).max() <= 1e-5 * float(g["cost_absmax"])
    assert np.abs(ow.reshape(-1)[::int(g["warped_stride"])] - g["warped_sub"]).max() <= 1e-5 * float(g["warped_absmax"])


CVP = cases.cvp_cases()


@pytest.mark.parametrize("c", CVP, ids=[c["name"] for c in CVP])
def test_cvp_oracle_matches_reference(c):
    """oracle.cvpmvsnet against the unmodified reference proj_cost / homo_warping (tools/make_cvp_golden.py)."""
    g = load_golden("cvp", c["name"])
    oc = oracle.cvpmvsnet.proj_cost(c["ref"], c["srcs"], c["ref_in"], c["src_in"], c["ref_ex"], c["src_ex"], c["depth_hypos"])
    ow = oracle.cvpmvsnet.homo_warping(c["srcs"][0], c["ref_in"], c["src_in"][:, 0], c["ref_ex"], c["src_ex"][:, 0],
                                       c["depth_planes"])
    assert np.abs(oc - g["cost"]).max() <= 2e-5 * np.abs(g["cost"]).max()
    assert np.abs(ow - g["warped"]).max() <= 2e-5 * np.abs(g["warped"]).max()


def test_fullsize_cfg4_oracle_matches_golden():
    """BASELINE.json configs[3] at full size (ETH3D 96x144 features, C=256, 256 planes, 2 views): the oracle against the golden of the
    unmodified reference (tools/make_fullsize_golden.py) -- masks, geometric masks and corner indices bit for bit."""
    case = cases.fullsize_planesweep_case()
    g = load_golden("ps", case["name"])
    corrs, masks, inv, dbg = oracle.planesweep_correlation(
        case["feat_key"], case["intrinsics_key"], case["feat_sources"], case["transforms"], case["intrinsics_sources"],
        return_debug=True, **case["kwargs"])
    check_planesweep_against_golden(g, corrs, masks, inv, [d["ix0"] for d in dbg], [d["iy0"] for d in dbg],
                                    [d["geom_mask"] for d in dbg], tol=1e-5)


def test_fullsize_cfg3_oracle_matches_golden():
    """BASELINE.json configs[2] at full size (DTU 224x304 features, C=32, 5 views, 192 planes): three 8-plane slabs of the oracle's
    variance volume against the strided subsample of the reference's full volume, and the regression at full size."""
    case = cases.fullsize_mvs_case()
    g = load_golden("mvs", case["name"])
    B, V, C, h, w = case["feats"].shape
    D = case["depth"].shape[1]
    sub, ref = int(g["sub"]), g["var_sub"]
    idx = np.arange(ref.size, dtype=np.int64) * sub                      # flat index into [B,C,D,h,w]
    d_of = (idx // (h * w)) % D
    checked = 0
    for d0 in (0, 96, 184):
        var = oracle.mvs_variance(case["feats"], case["proj"], case["depth"][:, d0:d0 + 8])      # [B,C,8,h,w]
        sel = (d_of >= d0) & (d_of < d0 + 8)
        c_of, p_of = idx[sel] // (D * h * w), idx[sel] % (h * w)
        got = var.reshape(C, 8, h * w)[c_of, d_of[sel] - d0, p_of]
        assert np.abs(got - ref[sel]).max() <= 1e-5 * float(g["var_absmax"])
        checked += int(sel.sum())
    assert checked > 40000
    d, c, i, e = oracle.mvs_regress(case["cost"], case["depth"])
    check_regress_against_golden(g, d, c, i, e, rtol=1e-5)


def test_torchcorr_border_padding_oracle_matches_golden():
    """TorchCorr(padding_mode="border") (planesweep_corr.py:143-148, :91-96): golden of the unmodified reference."""
    c = cases.border_case()
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "border_torchcorr.npz"))
    for nm in ("dim", False, True):
        for wm in (0, 1):
            corr, mask = oracle.torch_corr(c["feat_key"], c["feat_src"], c["grids"][:, :, 0], c["grids"][:, :, 1],
                                           mask=c["geom"] if wm else None, normalize=nm, padding_mode="border")
            assert np.array_equal(mask.astype(np.uint8), g[f"mask_{nm}_{wm}"])
            ref = g[f"corr_{nm}_{wm}"]
            assert np.abs(corr - ref).max() <= 1e-5 * np.abs(ref).max()


@pytest.mark.parametrize("case", cases.mvs_cases(include_large=False), ids=lambda c: c["name"])
def test_mvs_backward_oracle_matches_reference_autograd(case):
    """oracle.mvs_variance_backward / mvs_regress_backward against gradients frozen from the unmodified reference functions under
    CPU autograd (tools/make_mvs_bwd_golden.py)."""
    g = load_golden("mvsbwd", case["name"])
    g_var, g_depth = cases.mvs_upstream_grads(case)
    a = oracle.mvs_variance_backward(case["feats"], case["proj"], case["depth"], g_var)
    b = oracle.mvs_regress_backward(case["cost"], case["depth"], g_depth)
    assert np.abs(a - g["grad_feats"]).max() <= 2e-6 * np.abs(g["grad_feats"]).max()
    assert np.abs(b - g["grad_cost"]).max() <= 2e-6 * np.abs(g["grad_cost"]).max()
