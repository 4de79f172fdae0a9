"""Freeze golden vectors under tests/golden/ by running the UNMODIFIED reference code on CPU (container only).

Every .npz holds the inputs of one case of tools/cases.py plus what the reference returned: correlation volumes,
masks, integer sampling indices (replayed from the reference's own ``sampling_points[i].us/.vs``),
MVSNet variance volumes and regression outputs.  Large cases store packed masks, index digests and a strided
subsample of the floating-point volumes to keep the fixtures small.
"""
import hashlib
import os
import sys

import numpy as np
import torch
import torch.nn.functional as F

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(HERE))
import ref_loader  # noqa: E402
from tools import cases  # noqa: E402
from tools.validate_oracle import ref_indices  # noqa: E402

OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")
SUB = 31  # stride of the subsample kept for large volumes


def digest(a):
    return np.frombuffer(hashlib.sha256(np.ascontiguousarray(a).tobytes()).digest(), dtype=np.uint8)


def golden_planesweep(case):
    ps = ref_loader.planesweep_corr()
    t = lambda a: None if a is None else torch.from_numpy(np.asarray(a))  # noqa: E731
    kw = {k: (t(v) if isinstance(v, np.ndarray) else v) for k, v in case["kwargs"].items()}
    blk = ps.PlanesweepCorrelation(warp_only=case.get("warp_only", False), normalize=case.get("normalize", "dim"))
    corrs, masks, inv = blk(
        feat_key=t(case["feat_key"]), intrinsics_key=t(case["intrinsics_key"]),
        feat_sources=[t(f) for f in case["feat_sources"]], source_to_key_transforms=[t(x) for x in case["transforms"]],
        intrinsics_sources=None if case["intrinsics_sources"] is None else [t(x) for x in case["intrinsics_sources"]],
        **kw)
    large = corrs[0].numel() > 300_000
    out = dict(sampling_invdepths=inv.numpy(), large=np.array(large))
    for v in range(len(corrs)):
        sp = blk.sampling_points[v]
        hs, ws = case["feat_sources"][v].shape[-2:]
        ix0, iy0 = ref_indices(sp.us, sp.vs, hs, ws)
        c = corrs[v].numpy()
        m = masks[v].numpy()
        g = sp.mask.numpy()
        if large:
            out[f"corr_sub_{v}"] = c.reshape(-1)[::SUB].copy()
            out[f"corr_absmax_{v}"] = np.abs(c).max()
            out[f"mask_bits_{v}"] = np.packbits(m.astype(np.uint8))
            out[f"geom_bits_{v}"] = np.packbits(g.astype(np.uint8))
            out[f"ix0_sha_{v}"] = digest(ix0.astype(np.int32))
            out[f"iy0_sha_{v}"] = digest(iy0.astype(np.int32))
        else:
            out[f"corr_{v}"] = c
            out[f"mask_{v}"] = m.astype(np.uint8)
            out[f"geom_{v}"] = g.astype(np.uint8)
            out[f"ix0_{v}"] = ix0.astype(np.int16)
            out[f"iy0_{v}"] = iy0.astype(np.int16)
    return out


def golden_mvs(case):
    bu = ref_loader.block_utils()
    feats = torch.from_numpy(case["feats"])
    proj = torch.from_numpy(case["proj"])
    depth = torch.from_numpy(case["depth"])
    B, V, C, h, w = feats.shape
    D = depth.shape[1]
    ref_volume = feats[:, 0].unsqueeze(2).repeat(1, 1, D, 1, 1)
    vsum, vsq = ref_volume, ref_volume ** 2
    warped1 = None
    for v in range(1, V):
        wv = bu.homo_warp(feats[:, v], proj[:, v], proj[:, 0], depth)
        warped1 = wv if warped1 is None else warped1
        vsum = vsum + wv
        vsq = vsq + wv ** 2
    var = vsq.div_(V).sub_(vsum.div_(V).pow_(2)).numpy()
    cost = torch.from_numpy(case["cost"])
    prob = F.softmax(cost, 1)
    dep = bu.depth_regression(prob, depth)
    p4 = 4 * F.avg_pool3d(F.pad(prob.unsqueeze(1), pad=(0, 0, 0, 0, 1, 2)), (4, 1, 1), stride=1).squeeze(1)
    eidx = bu.depth_regression(prob, torch.arange(D, dtype=prob.dtype))
    didx = eidx.long()
    conf = torch.gather(p4, 1, didx.unsqueeze(1)).squeeze(1)
    pv, sa, pm = bu.soft_argmin(cost, dim=1, keepdim=True, window=2)
    ent = bu.entropy(pv, dim=1, keepdim=True)
    large = var.size > 300_000
    out = dict(large=np.array(large), depth_reg=dep.numpy(), conf=conf.numpy(), didx=didx.numpy().astype(np.int32),
               eidx=eidx.numpy(), soft_argmin=sa.numpy(), prob_map=pm.numpy(), entropy=ent.numpy())
    if large:
        out["var_sub"] = var.reshape(-1)[::SUB].copy()
        out["var_absmax"] = np.abs(var).max()
        out["warp1_sub"] = warped1.numpy().reshape(-1)[::SUB].copy()
    else:
        out["var"] = var
        out["warp1"] = warped1.numpy()
    return out


def main():
    os.makedirs(OUT, exist_ok=True)
    total = 0
    for c in cases.planesweep_cases():
        g = golden_planesweep(c)
        p = os.path.join(OUT, f"ps_{c['name']}.npz")
        np.savez_compressed(p, **g)
        total += os.path.getsize(p)
    for c in cases.mvs_cases():
        g = golden_mvs(c)
        p = os.path.join(OUT, f"mvs_{c['name']}.npz")
        np.savez_compressed(p, **g)
        total += os.path.getsize(p)
    print(f"wrote golden vectors to {OUT}: {total / 1e6:.2f} MB")


if __name__ == "__main__":
    main()
