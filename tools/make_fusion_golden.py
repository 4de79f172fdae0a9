"""Freeze golden vectors of the learned view fusion (container only: imports the UNMODIFIED reference module
/root/reference/rmvd/models/blocks/learned_fusion.py).  Per case: seeded module weights, corrs / masks -> the module's own
logits (its ``corr_to_view_weight``), ``view_weights``, ``fused_corr``, ``fused_mask``.  Writes tests/golden/fusion_*.npz.

    python tools/make_fusion_golden.py
"""
import importlib.util
import os

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "tests", "golden")
CASES = [dict(name="v2_small", V=2, N=1, h=6, w=10, seed=1), dict(name="v3_batch2", V=3, N=2, h=5, w=7, seed=2),
         dict(name="v5_allmasked", V=5, N=1, h=4, w=9, seed=3, zero_cols=True), dict(name="v1_passthrough", V=1, N=1, h=3, w=5, seed=4)]


def reference_module():
    spec = importlib.util.spec_from_file_location("_ref_learned_fusion", "/root/reference/rmvd/models/blocks/learned_fusion.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def make_inputs(c):
    rng = np.random.default_rng(c["seed"])
    S = 256  # the module's first convolution takes the planes as its 256 input channels
    corrs = [rng.standard_normal((c["N"], S, c["h"], c["w"])).astype(np.float32) * 3 for _ in range(c["V"])]
    masks = [(rng.random((c["N"], S, c["h"], c["w"])) < 0.6).astype(np.float32) for _ in range(c["V"])]
    if c.get("zero_cols"):
        for m in masks:
            m[:, :, :, :3] = 0.0      # pixels no view sees: fused mask 0, fused corr 0
    corrs = [cc * m for cc, m in zip(corrs, masks)]
    return corrs, masks


def main():
    mod = reference_module()
    for c in CASES:
        torch.manual_seed(c["seed"])
        blk = mod.LearnedFusion().eval()
        corrs, masks = make_inputs(c)
        tc, tm = [torch.from_numpy(x) for x in corrs], [torch.from_numpy(x) for x in masks]
        with torch.no_grad():
            logits = [blk.corr_to_view_weight(x).numpy() for x in tc]
            fc, fm = blk(tc, tm)
        out = dict(fused_corr=fc.numpy(), fused_mask=fm.numpy(), V=np.array(c["V"]))
        for v in range(c["V"]):
            out[f"logit_{v}"], out[f"corr_{v}"], out[f"mask_{v}"] = logits[v], corrs[v], masks[v].astype(np.uint8)
            if c["V"] > 1:
                out[f"weight_{v}"] = blk.view_weights[v].numpy()
        np.savez_compressed(os.path.join(OUT, f"fusion_{c['name']}.npz"), **out)
        print("wrote", c["name"], {k: v.shape for k, v in out.items() if k.startswith("fused")})


if __name__ == "__main__":
    main()
