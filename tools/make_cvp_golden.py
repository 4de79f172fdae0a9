"""Freeze golden vectors of CVP-MVSNet's ``proj_cost`` and ``homo_warping`` under
tests/golden/cvp_*.npz by running the UNMODIFIED ``/root/reference/rmvd/models/blocks/cvp_mvsnet_components.py`` functions
on CPU in this container; ``Tensor.cuda`` is stubbed to the identity for the run (both functions hard-code ``.cuda()``)."""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(HERE))
import ref_loader  # noqa: E402
from tools import cases  # noqa: E402

OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")


def main():
    import oracle
    cvp = ref_loader.cvp_components()
    t = torch.from_numpy
    real_cuda = torch.Tensor.cuda
    torch.Tensor.cuda = lambda self, *a, **k: self
    try:
        for c in cases.cvp_cases():
            settings = types.SimpleNamespace(nsrc=c["nsrc"], mode="train")
            src_feature = [[t(f)] for f in c["srcs"]]                         # src_feature[v][level]
            cost = cvp.proj_cost(settings, t(c["ref"]).clone(), src_feature, 0, t(c["ref_in"]), t(c["src_in"]), t(c["ref_ex"]),
                                 t(c["src_ex"]), t(c["depth_hypos"])).numpy()
            warped = cvp.homo_warping(t(c["srcs"][0]), t(c["ref_in"]), t(c["src_in"][:, 0]), t(c["ref_ex"]), t(c["src_ex"][:, 0]),
                                      t(c["depth_planes"])).numpy()
            np.savez_compressed(os.path.join(OUT, f"cvp_{c['name']}.npz"), cost=cost, warped=warped)
            oc = oracle.cvpmvsnet.proj_cost(c["ref"], c["srcs"], c["ref_in"], c["src_in"], c["ref_ex"], c["src_ex"], c["depth_hypos"])
            ow = oracle.cvpmvsnet.homo_warping(c["srcs"][0], c["ref_in"], c["src_in"][:, 0], c["ref_ex"], c["src_ex"][:, 0],
                                               c["depth_planes"])
            e1 = np.abs(oc - cost).max() / np.abs(cost).max()
            e2 = np.abs(ow - warped).max() / np.abs(warped).max()
            print(f"{c['name']:12s} cost {cost.shape} oracle vs reference: cost {e1:.2e} warped {e2:.2e} nonzero {np.mean(warped != 0):.2f}")
            assert e1 < 2e-5 and e2 < 2e-5
    finally:
        torch.Tensor.cuda = real_cuda


if __name__ == "__main__":
    main()
