"""Golden of TorchCorr(padding_mode="border") (planesweep_corr.py:143-195, constructor argument :144-148) from the UNMODIFIED
reference on CPU: caller grids that leave the source map on every side, with and without a geometric mask.
    python tools/make_border_golden.py  ->  tests/golden/border_torchcorr.npz"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(HERE))
import ref_loader  # noqa: E402
from tools import cases  # noqa: E402


def main():
    ps = ref_loader.planesweep_corr()
    c = cases.border_case()
    out = {}
    for nm in ("dim", False, True):
        for with_mask in (False, True):
            blk = ps.TorchCorr(normalize=nm, padding_mode="border")
            corr, mask = blk(torch.from_numpy(c["feat_key"]), torch.from_numpy(c["feat_src"]), torch.from_numpy(c["grids"]),
                             torch.from_numpy(c["geom"]) if with_mask else None)
            out[f"corr_{nm}_{int(with_mask)}"] = corr.numpy()
            out[f"mask_{nm}_{int(with_mask)}"] = mask.numpy().astype(np.uint8)
    p = os.path.join(os.path.dirname(HERE), "tests", "golden", "border_torchcorr.npz")
    np.savez_compressed(p, **out)
    print(p, os.path.getsize(p))


if __name__ == "__main__":
    main()
