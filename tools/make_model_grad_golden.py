"""Freeze golden parameter gradients of the UNMODIFIED reference ``robust_mvd`` model (route B, CPU, this container) for
the training direction of the hot path: seeded weights, ``input_adapter`` + ``forward`` with
autograd, ``loss = cases.model_grad_loss(aux)``, ``loss.backward()``.  The encoder gradients reach the encoder ONLY through
the correlation block's backward (and the context branch), so they pin the drop-in's backward inside the model.
Writes tests/golden/modelgrad_<case>.npz.  Run from the repo root:  python -m tools.make_model_grad_golden
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from robustmvd_b200 import synthetic as syn  # noqa: E402
from tools.ref_models import import_rmvd  # noqa: E402
from tools import cases  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")


def main():
    rmvd = import_rmvd()
    torch.set_num_threads(os.cpu_count())
    for name in cases.MODEL_GRAD_CASES:
        case = cases.MODEL_CASES[name]
        torch.manual_seed(cases.MODEL_SEED)
        model = rmvd.create_model(case["model"], pretrained=False, weights=None, train=False, num_gpus=0)
        sample = model.input_adapter(**syn.model_sample(**case["sample"]))
        pred, aux = model(**sample)
        loss = cases.model_grad_loss(aux)
        loss.backward()
        params = dict(model.named_parameters())
        out = dict(loss=np.float64(loss.item()))
        for pname in cases.MODEL_GRAD_PARAMS:
            g = params[pname].grad.numpy()
            out[pname] = g.reshape(-1)[::cases.MODEL_GRAD_STRIDE].copy()
            out[pname + "|absmax"] = np.abs(g).max()
        np.savez_compressed(os.path.join(GOLDEN, f"modelgrad_{name}.npz"), **out)
        print(name, float(loss), {k: float(np.abs(v).max()) for k, v in out.items() if "|" not in k})


if __name__ == "__main__":
    main()
