"""Pin ``oracle/`` against the UNMODIFIED reference code executed on CPU (container only).

Runs the reference ``PlanesweepCorrelation`` / ``homo_warp`` / MVSNet volume + regression on seeded inputs and
compares: masks and integer sampling indices bit for bit, floating-point outputs against max|ref|.
"""
import sys, os
import numpy as np
import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ref_loader  # noqa: E402
import oracle  # noqa: E402
from tools import cases  # noqa: E402


def ref_indices(us, vs, hs, ws):
    """SURVEY.md 8(c) index oracle: replay warp():87-88 and grid_sample's unnormalise with torch ops."""
    xg = 2 * us / ws - 1
    yg = 2 * vs / hs - 1
    # ATen CPU: fma(x + 1, size/2, -0.5) -- evaluate in float64 (exact product) then round once
    ix = ((xg + 1).double() * (ws / 2) - 0.5).float()
    iy = ((yg + 1).double() * (hs / 2) - 0.5).float()
    return (torch.floor(ix).clamp(-2, ws).int().numpy(), torch.floor(iy).clamp(-2, hs).int().numpy())


def check_planesweep(case, verbose=True):
    ps = ref_loader.planesweep_corr()
    t = lambda a: None if a is None else torch.from_numpy(np.asarray(a))  # noqa: E731
    kw = case["kwargs"]
    blk = ps.PlanesweepCorrelation(warp_only=case.get("warp_only", False), normalize=case.get("normalize", "dim"))
    tk = {k: (t(v) if isinstance(v, np.ndarray) else v) for k, v in kw.items()}
    corrs, masks, inv = blk(
        feat_key=t(case["feat_key"]), intrinsics_key=t(case["intrinsics_key"]),
        feat_sources=[t(f) for f in case["feat_sources"]],
        source_to_key_transforms=[t(x) for x in case["transforms"]],
        intrinsics_sources=None if case["intrinsics_sources"] is None else [t(x) for x in case["intrinsics_sources"]],
        **tk)
    ocorrs, omasks, oinv, dbg = oracle.planesweep_correlation(
        case["feat_key"], case["intrinsics_key"], case["feat_sources"], case["transforms"],
        case["intrinsics_sources"], warp_only=case.get("warp_only", False), normalize=case.get("normalize", "dim"),
        return_debug=True, **kw)
    res = dict(name=case["name"], inv_equal=bool(np.array_equal(inv.numpy(), oinv)))
    res["mask_mismatch"] = 0
    res["idx_mismatch"] = 0
    res["geom_mismatch"] = 0
    res["uv_mismatch"] = 0
    res["corr_relerr"] = 0.0
    for v in range(len(corrs)):
        sp = blk.sampling_points[v]
        hs, ws = case["feat_sources"][v].shape[-2:]
        rix, riy = ref_indices(sp.us, sp.vs, hs, ws)
        res["uv_mismatch"] += int((sp.us.numpy() != dbg[v]["us"]).sum() + (sp.vs.numpy() != dbg[v]["vs"]).sum())
        res["idx_mismatch"] += int((rix != dbg[v]["ix0"]).sum() + (riy != dbg[v]["iy0"]).sum())
        res["geom_mismatch"] += int((sp.mask.numpy() != dbg[v]["geom_mask"]).sum())
        res["mask_mismatch"] += int((masks[v].numpy() != omasks[v]).sum())
        r = corrs[v].numpy()
        res["corr_relerr"] = max(res["corr_relerr"], float(np.abs(r - ocorrs[v]).max() / max(np.abs(r).max(), 1e-30)))
        res["valid_frac"] = float(masks[v].numpy().mean())
    if verbose:
        print(res)
    return res


def check_mvs(case, verbose=True):
    bu = ref_loader.block_utils()
    feats = torch.from_numpy(case["feats"])
    proj = torch.from_numpy(case["proj"])
    depth = torch.from_numpy(case["depth"])
    B, V, C, h, w = feats.shape
    D = depth.shape[1]
    # mvsnet.py:124-135 replayed with the reference's homo_warp
    ref_volume = feats[:, 0].unsqueeze(2).repeat(1, 1, D, 1, 1)
    vsum = ref_volume
    vsq = ref_volume ** 2
    warped0 = None
    for v in range(1, V):
        wv = bu.homo_warp(feats[:, v], proj[:, v], proj[:, 0], depth)
        if warped0 is None:
            warped0 = wv
        vsum = vsum + wv
        vsq = vsq + wv ** 2
    var = vsq.div_(V).sub_(vsum.div_(V).pow_(2)).numpy()
    ovar = oracle.mvs_variance(case["feats"], case["proj"], case["depth"])
    ow, odbg = oracle.homo_warp(case["feats"][:, 1], case["proj"][:, 1], case["proj"][:, 0], case["depth"], return_debug=True)
    res = dict(name=case["name"])
    res["var_relerr"] = float(np.abs(var - ovar).max() / np.abs(var).max())
    res["warp_relerr"] = float(np.abs(warped0.numpy() - ow).max() / max(np.abs(warped0.numpy()).max(), 1e-30))
    res["warp_nonzero_frac"] = float((warped0.numpy() != 0).mean())
    # regression (mvsnet.py:140-158)
    cost = torch.from_numpy(case["cost"])
    prob = F.softmax(cost, 1)
    dep = bu.depth_regression(prob, depth)
    p4 = 4 * F.avg_pool3d(F.pad(prob.unsqueeze(1), pad=(0, 0, 0, 0, 1, 2)), (4, 1, 1), stride=1).squeeze(1)
    eidx = bu.depth_regression(prob, torch.arange(D, dtype=prob.dtype))
    didx = eidx.long()
    conf = torch.gather(p4, 1, didx.unsqueeze(1)).squeeze(1)
    od, oc, oi, oe = oracle.mvs_regress(case["cost"], case["depth"])
    res["depth_relerr"] = float((np.abs(dep.numpy() - od) / np.abs(dep.numpy())).max())
    same = didx.numpy() == oi
    res["idx_mismatch"] = int((~same).sum())
    res["conf_abserr"] = float(np.abs(conf.numpy() - oc)[same].max())
    if verbose:
        print(res)
    return res


if __name__ == "__main__":
    bad = 0
    for c in cases.planesweep_cases():
        r = check_planesweep(c)
        bad += r["mask_mismatch"] + r["idx_mismatch"] + r["geom_mismatch"] + r["uv_mismatch"] + (r["corr_relerr"] > 1e-5) + (not r["inv_equal"])
    for c in cases.mvs_cases():
        r = check_mvs(c)
        bad += (r["var_relerr"] > 1e-5) + (r["depth_relerr"] > 1e-5)
    print("ORACLE PINNED" if bad == 0 else f"ORACLE MISMATCHES: {bad}")
