"""Freeze golden gradients of the MVSNet cost-volume path under tests/golden/mvsbwd_*.npz by running the UNMODIFIED reference
functions on CPU with autograd (container only): ``homo_warp`` (rmvd/models/blocks/utils.py:222-268) inside the volume assembly of
rmvd/models/mvsnet.py:123-136, and ``F.softmax`` + ``depth_regression`` (mvsnet.py:140-141, utils.py:271-274).

Per case of ``tools/cases.py:mvs_cases`` (small ones): grad of ``<g_var, variance>`` with respect to ``feats`` and grad of
``<g_depth, depth>`` with respect to the regularised cost; the upstream gradients are regenerated from the case's seed
(``tools/cases.py:mvs_upstream_grads``)."""
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

OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")


def reference_grads(case):
    bu = ref_loader.block_utils()
    feats = torch.from_numpy(case["feats"]).clone().requires_grad_(True)
    proj, depth = torch.from_numpy(case["proj"]), torch.from_numpy(case["depth"])
    B, V, C, h, w = feats.shape
    D = depth.shape[1]
    ref_volume = feats[:, 0].unsqueeze(2).repeat(1, 1, D, 1, 1)          # mvsnet.py:123-126
    vsum, vsq = ref_volume, ref_volume ** 2
    for v in range(1, V):                                                # :128-133
        wv = bu.homo_warp(feats[:, v], proj[:, v], proj[:, 0], depth)
        vsum = vsum + wv
        vsq = vsq + wv ** 2
    var = vsq.div_(V).sub_(vsum.div_(V).pow_(2))                         # :135
    g_var, g_depth = cases.mvs_upstream_grads(case)
    (g_feats,) = torch.autograd.grad((var * torch.from_numpy(g_var)).sum(), [feats])
    cost = torch.from_numpy(case["cost"]).clone().requires_grad_(True)
    dep = bu.depth_regression(F.softmax(cost, 1), depth)                 # :140-141
    (g_cost,) = torch.autograd.grad((dep * torch.from_numpy(g_depth)).sum(), [cost])
    return g_feats.numpy(), g_cost.numpy()


def main():
    os.makedirs(OUT, exist_ok=True)
    for case in cases.mvs_cases(include_large=False):
        g_feats, g_cost = reference_grads(case)
        np.savez_compressed(os.path.join(OUT, f"mvsbwd_{case['name']}.npz"), grad_feats=g_feats, grad_cost=g_cost)
        print(f"{case['name']:16s} |grad_feats| max {np.abs(g_feats).max():.3e}  |grad_cost| max {np.abs(g_cost).max():.3e}")


if __name__ == "__main__":
    main()
