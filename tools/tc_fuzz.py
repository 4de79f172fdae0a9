"""Randomised parity run of the plane-sweep correlation (tensor-core kernel path) against the numpy oracle: random map sizes,
channel counts, poses, plane counts (1 .. > 256: plane-table rebuilds) and explicit inverse-depth schedules that are
non-monotone and contain zero / negative / inf / nan entries.  Masks and integer indices must be bit-exact, corr within 1e-4
of max|corr|.  Every case runs with and without the debug outputs (the latter enables dead-plane skipping).

    python tools/tc_fuzz.py [cases] [seed]       # prints one line per case, exits 1 on the first mismatch
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oracle  # noqa: E402
from robustmvd_b200 import synthetic as syn  # noqa: E402
from robustmvd_b200.blocks.planesweep_corr import PlanesweepCorrelation  # noqa: E402

TOL = 1e-4


def make_case(rng):
    h, w = int(rng.integers(1, 41)), int(rng.integers(1, 101))
    same = rng.random() < 0.6
    hs, ws = (h, w) if same else (int(rng.integers(1, 50)), int(rng.integers(1, 120)))
    C = int(rng.choice([32, 64, 96]))
    V = int(rng.integers(1, 4))
    n = int(rng.integers(1, 3))
    S = int(rng.choice([1, 2, 3, 17, 40, 64, 130, 257, 300]))
    feat_key = syn.features((n, C, h, w), int(rng.integers(1 << 30)))
    feat_sources = [syn.features((n, C, hs, ws), int(rng.integers(1 << 30))) for _ in range(V)]
    K = syn.relative_intrinsics(h, w, fx_rel=float(rng.uniform(0.3, 1.2)), n=n)
    Ts = []
    for _ in range(V):
        big = rng.random() < 0.3
        a = 40.0 if big else 6.0
        R = syn.rot_y(rng.uniform(-a, a)) @ syn.rot_x(rng.uniform(-a / 2, a / 2)) @ syn.rot_z(rng.uniform(-a, a))
        t = tuple(rng.uniform(-0.8, 0.8, 3) * (1.0 if rng.random() < 0.8 else 0.0))
        Ts.append(syn.make_transform(R, t, n=n))
    kw = {}
    if rng.random() < 0.5:
        kw = dict(num_sampling_points=S, min_depth=float(rng.uniform(0.2, 2.0)), max_depth=float(rng.uniform(5.0, 1000.0)),
                  sampling_type=str(rng.choice(["linear_invdepth", "linear_depth"])))
    else:
        inv = rng.uniform(0.001, 3.0, (n, S)).astype(np.float32)
        if rng.random() < 0.5:
            inv.sort(axis=1)
        for special in (0.0, -0.5, np.inf, -np.inf, np.nan, 1e-30, 1e30):
            if rng.random() < 0.25:
                inv[int(rng.integers(n)), int(rng.integers(S))] = special
        kw = dict(sampling_invdepths=inv)
    return dict(feat_key=feat_key, intrinsics_key=K, feat_sources=feat_sources, transforms=Ts, kwargs=kw,
                desc=f"h{h} w{w} hs{hs} ws{ws} C{C} V{V} n{n} S{S} {'range' if 'min_depth' in kw else 'explicit'}")


def run_cuda(case, debug):
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
    kw = {k: (t(v) if isinstance(v, np.ndarray) else v) for k, v in case["kwargs"].items()}
    blk = PlanesweepCorrelation()
    corrs, masks, _ = blk(feat_key=t(case["feat_key"]), intrinsics_key=t(case["intrinsics_key"]),
                          feat_sources=[t(f) for f in case["feat_sources"]],
                          source_to_key_transforms=[t(x) for x in case["transforms"]], debug=debug, **kw)
    torch.cuda.synchronize()
    dbg = blk.debug
    return ([c.cpu().numpy() for c in corrs], [m.cpu().numpy() for m in masks],
            None if dbg is None else {k: v.cpu().numpy() for k, v in dbg.items()})


def main():
    ncases = int(sys.argv[1]) if len(sys.argv) > 1 else 60
    rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 7)
    for i in range(ncases):
        case = make_case(rng)
        with np.errstate(all="ignore"):
            oc, om, _, od = oracle.planesweep_correlation(case["feat_key"], case["intrinsics_key"], case["feat_sources"],
                                                          case["transforms"], None, return_debug=True, **case["kwargs"])
        corrs, masks, dbg = run_cuda(case, True)
        corrs2, masks2, _ = run_cuda(case, False)
        worst = 0.0
        for v in range(len(corrs)):
            ok = np.array_equal(masks[v], om[v]) and np.array_equal(dbg["ix0"][v], od[v]["ix0"]) and \
                np.array_equal(dbg["iy0"][v], od[v]["iy0"]) and np.array_equal(dbg["geom"][v].astype(bool), od[v]["geom_mask"])
            ok = ok and np.array_equal(masks2[v], masks[v]) and np.array_equal(corrs2[v], corrs[v], equal_nan=True)
            scale = max(float(np.abs(oc[v]).max()), 1e-30)
            err = float(np.abs(corrs[v] - oc[v]).max()) / scale
            worst = max(worst, err)
            if not ok or not (err <= TOL):
                print(f"case {i} MISMATCH view {v}: {case['desc']} err {err} masks {int((masks[v] != om[v]).sum())}", flush=True)
                return 1
        print(f"case {i} ok  {case['desc']}  valid {float(np.mean([m.mean() for m in masks])):.3f} err {worst:.2e}", flush=True)
    print("fuzz done", ncases, "cases, 0 mismatches")
    return 0


if __name__ == "__main__":
    sys.exit(main())
