"""Helpers shared by the CPU (oracle vs golden) and GPU (CUDA vs oracle vs golden) parity tests."""
import hashlib
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
SUB = 31  # stride of the subsample kept for large volumes (tools/make_golden.py)


def load_golden(prefix, name):
    return np.load(os.path.join(GOLDEN, f"{prefix}_{name}.npz"))


def digest(a):
    return np.frombuffer(hashlib.sha256(np.ascontiguousarray(a).tobytes()).digest(), dtype=np.uint8)


def check_planesweep_against_golden(g, corrs, masks, inv, ix0s=None, iy0s=None, geoms=None, tol=1e-4):
    """corrs/masks: lists of numpy arrays per view.  Masks / indices bit-exact, corr within tol * max|ref|."""
    assert np.array_equal(np.asarray(inv, np.float32).reshape(g["sampling_invdepths"].shape), g["sampling_invdepths"])
    large = bool(g["large"])
    for v in range(len(corrs)):
        c = np.asarray(corrs[v], np.float32)
        m = np.asarray(masks[v], np.float32)
        if large:
            assert np.array_equal(np.packbits(m.astype(np.uint8)), g[f"mask_bits_{v}"]), f"mask view {v}"
            ref = g[f"corr_sub_{v}"]
            scale = max(float(g[f"corr_absmax_{v}"]), 1e-30)
            assert np.abs(c.reshape(-1)[::SUB] - ref).max() <= tol * scale, f"corr view {v}"
            if ix0s is not None:
                assert np.array_equal(digest(ix0s[v].astype(np.int32)), g[f"ix0_sha_{v}"]), f"ix0 view {v}"
                assert np.array_equal(digest(iy0s[v].astype(np.int32)), g[f"iy0_sha_{v}"]), f"iy0 view {v}"
            if geoms is not None:
                assert np.array_equal(np.packbits(geoms[v].astype(np.uint8)), g[f"geom_bits_{v}"]), f"geom view {v}"
        else:
            assert np.array_equal(m.astype(np.uint8), g[f"mask_{v}"]), f"mask view {v}"
            ref = g[f"corr_{v}"]
            scale = max(float(np.abs(ref).max()), 1e-30)
            assert np.abs(c - ref).max() <= tol * scale, f"corr view {v}: {np.abs(c - ref).max() / scale}"
            if ix0s is not None:
                assert np.array_equal(ix0s[v].astype(np.int16), g[f"ix0_{v}"]), f"ix0 view {v}"
                assert np.array_equal(iy0s[v].astype(np.int16), g[f"iy0_{v}"]), f"iy0 view {v}"
            if geoms is not None:
                assert np.array_equal(geoms[v].astype(np.uint8), g[f"geom_{v}"]), f"geom view {v}"


def check_mvs_variance_against_golden(g, var, tol=1e-4):
    if bool(g["large"]):
        ref = g["var_sub"]
        sub = int(g["sub"]) if "sub" in g.files else SUB      # full-size fixtures keep a coarser subsample
        assert np.abs(var.reshape(-1)[::sub] - ref).max() <= tol * float(g["var_absmax"])
    else:
        ref = g["var"]
        assert np.abs(var - ref).max() <= tol * float(np.abs(ref).max())


def check_warp_against_golden(g, warped, tol=1e-4):
    if bool(g["large"]):
        ref = g["warp1_sub"]
        assert np.abs(warped.reshape(-1)[::SUB] - ref).max() <= tol * max(float(np.abs(ref).max()), 1e-30)
    else:
        ref = g["warp1"]
        assert np.abs(warped - ref).max() <= tol * max(float(np.abs(ref).max()), 1e-30)


def check_regress_against_golden(g, depth, conf, didx, eidx, rtol=1e-3):
    """depth within rtol relative; index exact except where the expected index sits within 1e-4 of an integer."""
    assert (np.abs(depth - g["depth_reg"]) <= rtol * np.abs(g["depth_reg"])).all()
    near = np.abs(g["eidx"] - np.round(g["eidx"])) < 1e-3
    same = didx.astype(np.int64) == g["didx"].astype(np.int64)
    assert (same | near).all()
    assert np.abs(conf - g["conf"])[same].max() <= 1e-4
    assert np.abs(eidx - g["eidx"]).max() <= 1e-3 * max(1.0, float(np.abs(g["eidx"]).max()))
