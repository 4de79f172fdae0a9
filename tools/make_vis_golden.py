"""Freeze golden vectors of the Vis-MVSNet cost volume under tests/golden/vis_*.npz by running the
UNMODIFIED reference functions ``get_homographies``, ``homography_warping`` and ``groupwise_correlation``
(``/root/reference/rmvd/models/blocks/utils.py:71-189``) on CPU in this container, composed exactly as
``vis_mvsnet_singlestage.py:86-122, :230-242`` composes them.  ``get_pixel_grids`` (:155-160) hard-codes ``.cuda()``: the
harness stubs ``Tensor.cuda`` to the identity for the duration of the run (no reference file is modified)."""
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


def reference_cost_volume(case):
    bu = ref_loader.block_utils()
    t = torch.from_numpy
    ref, src = t(case["ref"]), t(case["src"])
    d = case["depth_num"]
    Hs = bu.get_homographies(t(case["left_cam"]), t(case["right_cam"]), d, t(case["depth_start"]), t(case["depth_interval"]),
                             inv=case["inv"])                                                          # ndhw33
    src_nd = src.unsqueeze(1).repeat(1, d, 1, 1, 1).view(-1, *src.size()[1:])                          # :109-113
    warped = bu.homography_warping(src_nd, Hs.view(-1, *Hs.size()[2:]))                                # :114-116
    warped = warped.view(-1, d, *src.size()[1:]).transpose(1, 2)                                       # ncdhw, :117-121
    ref_ncdhw = ref.unsqueeze(2).repeat(1, 1, d, 1, 1)
    cost = bu.groupwise_correlation(ref_ncdhw, warped, case["groups"], 1)                              # :242
    return Hs.numpy(), warped.contiguous().numpy(), cost.numpy()


def main():
    import oracle
    real_cuda = torch.Tensor.cuda
    torch.Tensor.cuda = lambda self, *a, **k: self
    try:
        for case in cases.vis_cases():
            Hs, warped, cost = reference_cost_volume(case)
            big = cost.size > 300_000   # large volumes: strided subsample
            np.savez_compressed(os.path.join(OUT, f"vis_{case['name']}.npz"), Hs=Hs,
                                cost_sub=cost.reshape(-1)[::(7 if big else 1)].copy(), cost_stride=np.int64(7 if big else 1),
                                cost_absmax=np.abs(cost).max(), warped_sub=warped.reshape(-1)[::(31 if big else 7)].copy(),
                                warped_stride=np.int64(31 if big else 7), warped_absmax=np.abs(warped).max())
            ow = oracle.homography_warp_volume(case["src"], Hs)
            oc = oracle.groupwise_cost_volume(case["ref"], case["src"], Hs, case["groups"])
            e1 = np.abs(ow - warped).max() / np.abs(warped).max()
            e2 = np.abs(oc - cost).max() / np.abs(cost).max()
            print(f"{case['name']:16s} Hs {Hs.shape} cost {cost.shape}  oracle vs reference: warped {e1:.2e} cost {e2:.2e}  "
                  f"nonzero {np.mean(warped != 0):.2f}")
            assert e1 < 1e-5 and e2 < 1e-5
    finally:
        torch.Tensor.cuda = real_cuda


if __name__ == "__main__":
    main()
