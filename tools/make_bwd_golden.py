"""Freeze golden gradients of the correlation block under tests/golden/psbwd_*.npz by running the UNMODIFIED reference
``PlanesweepCorrelation`` on CPU with autograd (container only).

Each file holds what ``torch.autograd.grad(sum_v <grad_corrs[v], corrs[v]>, [feat_key, *feat_sources])`` returned for
one case of ``tools/cases.py:backward_cases`` (inputs and upstream gradients are regenerated from the seeds)."""
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
    t = lambda a: None if a is None else torch.from_numpy(np.asarray(a))  # noqa: E731
    kw = {k: (t(v) if isinstance(v, np.ndarray) else v) for k, v in case["kwargs"].items()}
    key = t(case["feat_key"]).clone().requires_grad_(True)
    srcs = [t(f).clone().requires_grad_(True) for f in case["feat_sources"]]
    blk = ps.PlanesweepCorrelation(warp_only=False, normalize=case.get("normalize", "dim"))
    corrs, masks, inv = blk(
        feat_key=key, intrinsics_key=t(case["intrinsics_key"]), feat_sources=srcs,
        source_to_key_transforms=[t(x) for x in case["transforms"]],
        intrinsics_sources=None if case["intrinsics_sources"] is None else [t(x) for x in case["intrinsics_sources"]],
        **kw)
    S = corrs[0].shape[1]
    gs = cases.grad_corrs(case, S)
    loss = sum((c * t(g)).sum() for c, g in zip(corrs, gs))
    grads = torch.autograd.grad(loss, [key] + srcs)
    return [g.numpy() for g in grads], [m.numpy() for m in masks]


def main():
    import oracle
    os.makedirs(OUT, exist_ok=True)
    worst = 0.0
    for case in cases.backward_cases():
        grads, masks = reference_backward(case)
        out = dict(grad_key=grads[0])
        for v, g in enumerate(grads[1:]):
            out[f"grad_src_{v}"] = g
        np.savez_compressed(os.path.join(OUT, f"psbwd_{case['name']}.npz"), **out)
        # pin the oracle restatement against the reference's autograd in the same pass
        S = masks[0].shape[1]
        ogk, ogs = oracle.planesweep_correlation_backward(
            case["feat_key"], case["intrinsics_key"], case["feat_sources"], case["transforms"],
            cases.grad_corrs(case, S), case["intrinsics_sources"], normalize=case.get("normalize", "dim"), **case["kwargs"])
        errs = [np.abs(ogk - grads[0]).max() / max(np.abs(grads[0]).max(), 1e-30)]
        errs += [np.abs(a - b).max() / max(np.abs(b).max(), 1e-30) for a, b in zip(ogs, grads[1:])]
        worst = max(worst, max(errs))
        print(f"{case['name']:28s} oracle vs reference autograd: max rel err {max(errs):.2e}")
    print("worst", worst)
    assert worst < 1e-5


if __name__ == "__main__":
    main()
