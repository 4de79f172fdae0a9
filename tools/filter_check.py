"""CPU model of the filtered locate of csrc/psweep_ws.cu (sample_fast): for every sample the cheap chain settles,
its integer corner and its mask must equal the oracle's exact chain.  The reciprocal is modelled as the correctly
rounded 1/den perturbed by up to +-2 ulp (rcp.approx.ftz.f32 is specified to 1 ulp) before the Newton step.

    python tools/filter_check.py        # prints settled fraction and mismatch counts per case (must be 0)
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import planesweep as op  # noqa: E402
from robustmvd_b200 import synthetic as syn  # noqa: E402

F = np.float32


def fma32(a, b, c):
    return (a.astype(np.float64) * b.astype(np.float64) + np.float64(c)).astype(np.float32) if np.isscalar(c) else \
        (a.astype(np.float64) * b.astype(np.float64) + c.astype(np.float64)).astype(np.float32)


def fast_chain(coeffs, invd, hs, ws, rng, ulps=2):
    u_h, v_h, k_h, m_u, m_v, m_k = coeffs
    ds = invd.astype(np.float32)
    with np.errstate(all="ignore"):
        den = (k_h + m_k * ds).astype(np.float32)
        nu = (u_h + m_u * ds).astype(np.float32)
        nv = (v_h + m_v * ds).astype(np.float32)
        r0 = (F(1) / den).astype(np.float32)
        r0 = (r0 * (F(1) + rng.integers(-ulps, ulps + 1, den.shape).astype(np.float32) * F(2.0 ** -23))).astype(np.float32)
        r = fma32(r0, fma32(-den, r0, F(1)), r0)
        ix = fma32(nu, r, F(-0.5))
        iy = fma32(nv, r, F(-0.5))
        fx = ix - np.floor(ix)
        fy = iy - np.floor(iy)
        xi = np.where(np.isfinite(ix), np.floor(np.clip(ix, -2e9, 2e9)), 0).astype(np.int64)
        yi = np.where(np.isfinite(iy), np.floor(np.clip(iy, -2e9, 2e9)), 0).astype(np.int64)
        kU = F(2.0 ** -24)
        eps_x, eps_y = kU * (F(16) * F(ws) + F(64)), kU * (F(16) * F(hs) + F(64))
        int_x = (xi >= 0) & (xi < ws - 1)
        int_y = (yi >= 0) & (yi < hs - 1)
        tx = np.minimum(fx, F(1) - fx) > np.where(int_x, eps_x, eps_x + F(1.2e-4))
        ty = np.minimum(fy, F(1) - fy) > np.where(int_y, eps_y, eps_y + F(1.2e-4))
        den_ok = np.abs(den) <= F(2.0 ** 60)
    settled = den_ok & tx & ty
    return settled, np.clip(xi, -2, ws), np.clip(yi, -2, hs), int_x & int_y


def check(name, h, w, hs, ws, Kk, Ks, T, invd, rng):
    coeffs = op.epipolar_coeffs(Kk, Ks, T, h, w, hs, ws)
    us, vs, zs, geom = op.sampling_points(coeffs, invd)
    b = op.bilinear_setup(us, vs, hs, ws)
    mask = (b["samp_mask"] > 0) & geom
    settled, x0, y0, interior = fast_chain(coeffs, invd, hs, ws, rng)
    settled = np.broadcast_to(settled, us.shape)
    bad_x = int((settled & (x0 != b["ix0"])).sum())
    bad_y = int((settled & (y0 != b["iy0"])).sum())
    bad_m = int((settled & ((interior & geom) != mask)).sum())
    print(f"{name:28s} samples {us.size:9d} settled {settled.mean():.5f} valid {mask.mean():.3f} "
          f"bad_x {bad_x} bad_y {bad_y} bad_mask {bad_m}")
    return bad_x + bad_y + bad_m


def main():
    rng = np.random.default_rng(0)
    bad = 0
    for cfg, V in ((syn.CFG2, 2), (syn.CFG4, 4)):
        h, w = cfg["h"], cfg["w"]
        inp = syn.planesweep_inputs(h, w, 4, V)
        invd = op.compute_sampling_invdepths(cfg["min_depth"], cfg["max_depth"], cfg["S"])[:, :, None, None]
        for v in range(V):
            bad += check(f"cfg {h}x{w} view {v}", h, w, h, w, inp["intrinsics_key"], inp["intrinsics_key"],
                         inp["transforms"][v], invd, rng)
    # stress: random poses, source resolution != key resolution
    for k in range(12):
        h, w, hs, ws = 24, 40, int(rng.integers(9, 60)), int(rng.integers(9, 80))
        R = syn.rot_y(rng.uniform(-20, 20)) @ syn.rot_x(rng.uniform(-10, 10)) @ syn.rot_z(rng.uniform(-15, 15))
        T = syn.make_transform(R, tuple(rng.uniform(-0.6, 0.6, 3)))
        invd = op.compute_sampling_invdepths(0.4, 1000.0, 128)[:, :, None, None]
        K = syn.relative_intrinsics(h, w)
        bad += check(f"random {k} src {hs}x{ws}", h, w, hs, ws, K, K, T, invd, rng)
    print("TOTAL MISMATCHES", bad)
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
