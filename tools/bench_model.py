"""Device time of one robust_mvd forward (random-init weights, synthetic KITTI-size sample: 384x1280, V source views).

    python tools/bench_model.py [--views 2] [--reps 10]        # RMVD_BATCH_ENCODER=0 for the per-view encoder calls
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import robustmvd_b200 as rmvd  # noqa: E402
from robustmvd_b200 import synthetic as syn  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--views", type=int, default=2)
    ap.add_argument("--reps", type=int, default=10)
    args = ap.parse_args()
    torch.manual_seed(0)
    model = rmvd.create_model("robust_mvd", pretrained=False, weights=None, train=False, num_gpus=1)
    sample = syn.model_sample(384, 1280, args.views, batch=1)
    inp = model.input_adapter(**sample)
    with torch.no_grad():
        for _ in range(3):
            model(**inp)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.reps):
            model(**inp)
        e1.record()
        torch.cuda.synchronize()
    print(json.dumps({"model": "robust_mvd", "image": "384x1280", "source_views": args.views,
                      "batched_encoder": os.environ.get("RMVD_BATCH_ENCODER", "1") != "0",
                      "ms_per_forward": e0.elapsed_time(e1) / args.reps}))


if __name__ == "__main__":
    main()
