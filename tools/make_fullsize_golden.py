"""Full-size goldens of BASELINE.json configs[2] (MVSNet, DTU 224x304 features, C=32, 5 views, 192 planes) and configs[3]
(robust_mvd, ETH3D 96x144 features, C=256, 256 planes, 2 source views) from the UNMODIFIED reference on CPU (container only),
in the packed "large" format of tools/make_golden.py: packed masks, index digests and a strided subsample of the float volumes.

    python tools/make_fullsize_golden.py        # ~3 min, ~8 GB of host memory
"""
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
from tools.make_golden import golden_planesweep, OUT  # noqa: E402

SUB_VAR = 997    # stride of the subsample of the 418 M-element variance volume (prime: walks every channel / plane / row phase)


def golden_cfg3(case, slab=16):
    bu = ref_loader.block_utils()
    feats, proj, depth = (torch.from_numpy(case[k]) for k in ("feats", "proj", "depth"))
    B, V, C, h, w = feats.shape
    D = depth.shape[1]
    var = np.empty((B, C, D, h, w), np.float32)
    for d0 in range(0, D, slab):                     # mvsnet.py:123-135 on a range of planes (the planes are independent)
        dv = depth[:, d0:d0 + slab]
        ref_volume = feats[:, 0].unsqueeze(2).repeat(1, 1, dv.shape[1], 1, 1)
        vsum, vsq = ref_volume, ref_volume ** 2
        for v in range(1, V):
            wv = bu.homo_warp(feats[:, v], proj[:, v], proj[:, 0], dv)
            vsum = vsum + wv
            vsq = vsq + wv ** 2
        var[:, :, d0:d0 + slab] = vsq.div_(V).sub_(vsum.div_(V).pow_(2)).numpy()
    cost = torch.from_numpy(case["cost"])
    prob = F.softmax(cost, 1)                                                                    # mvsnet.py:140
    dep = bu.depth_regression(prob, depth)                                                       # :141
    p4 = 4 * F.avg_pool3d(F.pad(prob.unsqueeze(1), pad=(0, 0, 0, 0, 1, 2)), (4, 1, 1), stride=1).squeeze(1)   # :146-148
    eidx = bu.depth_regression(prob, torch.arange(D, dtype=prob.dtype))
    didx = eidx.long()
    conf = torch.gather(p4, 1, didx.unsqueeze(1)).squeeze(1)
    return dict(large=np.array(True), sub=np.array(SUB_VAR), var_sub=var.reshape(-1)[::SUB_VAR].copy(), var_absmax=np.abs(var).max(),
                depth_reg=dep.numpy(), conf=conf.numpy(), didx=didx.numpy().astype(np.int32), eidx=eidx.numpy())


def main():
    c4 = cases.fullsize_planesweep_case()
    g = golden_planesweep(c4)
    assert bool(g["large"])
    p = os.path.join(OUT, f"ps_{c4['name']}.npz")
    np.savez_compressed(p, **g)
    print(p, os.path.getsize(p) / 1e6, "MB")
    c3 = cases.fullsize_mvs_case()
    g = golden_cfg3(c3)
    p = os.path.join(OUT, f"mvs_{c3['name']}.npz")
    np.savez_compressed(p, **g)
    print(p, os.path.getsize(p) / 1e6, "MB")


if __name__ == "__main__":
    main()
