"""Freeze golden gradients of the warp-only block under tests/golden/warpbwd_*.npz by running the UNMODIFIED reference
``PlanesweepCorrelation(warp_only=True, normalize=...)`` (rmvd/models/blocks/planesweep_corr.py:107-140, :396-427) on CPU with autograd
(container only).  Per case of ``tools/cases.py:warp_backward_cases``: ``torch.autograd.grad(sum_v <g_v, warped_v>, feat_sources)`` for
seeded upstream gradients ``g_v`` (``cases.warp_upstream_grads``)."""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(HERE))
import ref_loader  # noqa: E402
from tools import cases  # noqa: E402

OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")


def reference_backward(case):
    ps = ref_loader.planesweep_corr()
    t = lambda a: torch.from_numpy(np.asarray(a))  # noqa: E731
    srcs = [t(f).clone().requires_grad_(True) for f in case["feat_sources"]]
    blk = ps.PlanesweepCorrelation(warp_only=True, normalize=case["normalize"])
    warped, masks, _ = blk(feat_key=t(case["feat_key"]), intrinsics_key=t(case["intrinsics_key"]), feat_sources=srcs,
                           source_to_key_transforms=[t(x) for x in case["transforms"]], **case["kwargs"])
    gs = cases.warp_upstream_grads(case, warped[0].shape)
    loss = sum((wv * t(g)).sum() for wv, g in zip(warped, gs))
    return [g.numpy() for g in torch.autograd.grad(loss, srcs)]


def main():
    os.makedirs(OUT, exist_ok=True)
    for case in cases.warp_backward_cases():
        grads = reference_backward(case)
        np.savez_compressed(os.path.join(OUT, f"warpbwd_{case['name']}.npz"), **{f"grad_src_{v}": g for v, g in enumerate(grads)})
        print(f"{case['name']:24s} max|grad| {max(np.abs(g).max() for g in grads):.3e}")


if __name__ == "__main__":
    main()
