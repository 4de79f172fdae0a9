"""Freeze model-level golden vectors from the UNMODIFIED reference models run on CPU in this container
(route B, tools/ref_models.py): ``rmvd.create_model(name, pretrained=False, num_gpus=0)`` with
``torch.manual_seed(SEED)`` before construction, ``model.run(**sample)`` on ``synthetic.model_sample``.

Writes tests/golden/model_<name>.npz (predictions) and tests/golden/model_<name>_state.json (parameter names, shapes
and float64 sums of the seeded initialisation: the candidate must reproduce them to share the weights).
Run from the repo root:  python -m tools.make_model_golden
"""
import json
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


def state_summary(model):
    return {k: dict(shape=list(v.shape), sum=float(v.double().sum()), abssum=float(v.double().abs().sum()))
            for k, v in model.state_dict().items()}


def main():
    rmvd = import_rmvd()
    torch.set_num_threads(os.cpu_count())
    for name, case in cases.MODEL_CASES.items():
        model_name = case["model"]
        torch.manual_seed(cases.MODEL_SEED)
        model = rmvd.create_model(model_name, pretrained=False, weights=None, train=False, num_gpus=0)
        with open(os.path.join(GOLDEN, f"model_{model_name}_state.json"), "w") as f:
            json.dump(state_summary(model), f, indent=0, sort_keys=True)
        cases.prepare_model_for_case(model, model_name)
        sample = syn.model_sample(**case["sample"])
        pred, aux = model.run(**sample)
        out = {f"pred_{k}": np.asarray(v) for k, v in pred.items()}
        for k in ("invdepth", "invdepth_log_b"):
            if k in aux:
                out[f"aux_{k}"] = np.asarray(aux[k])
        np.savez_compressed(os.path.join(GOLDEN, f"model_{name}.npz"), **out)
        print(name, {k: (v.shape, float(np.nanmin(v)), float(np.nanmax(v))) for k, v in out.items()})


if __name__ == "__main__":
    main()
