"""Where the wall time of ``model.run(**sample)`` goes (robust_mvd, random-init weights, synthetic KITTI-size sample): batching of the
numpy inputs, input_adapter (normalisation + host-to-device copies), forward (host launch time and device time), output_adapter
(device-to-host copies).  One JSON line.

    python tools/profile_model_run.py [--views 2] [--reps 10]
"""
import argparse
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import robustmvd_b200 as rmvd  # noqa: E402
from robustmvd_b200 import synthetic as syn  # noqa: E402
from robustmvd_b200.models.helpers import add_batch_dim  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--views", type=int, default=2)
    ap.add_argument("--reps", type=int, default=10)
    args = ap.parse_args()
    torch.manual_seed(0)
    model = rmvd.create_model("robust_mvd", pretrained=False, weights=None, train=False, num_gpus=1)
    sample = syn.model_sample(384, 1280, args.views)
    for _ in range(3):
        model.run(**sample)
    torch.cuda.synchronize()
    acc = {"batch": 0.0, "input_adapter": 0.0, "forward_launch": 0.0, "forward_wait": 0.0, "output_adapter": 0.0, "run_total": 0.0}
    dev_ms = 0.0
    with torch.no_grad():
        for _ in range(args.reps):
            t0 = time.perf_counter()
            b = add_batch_dim(sample["images"], sample["keyview_idx"], sample.get("poses"), sample.get("intrinsics"), sample.get("depth_range"))
            t1 = time.perf_counter()
            inp = model.input_adapter(images=b[0], keyview_idx=b[1], poses=b[2], intrinsics=b[3], depth_range=b[4])
            torch.cuda.synchronize()
            t2 = time.perf_counter()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            out = model(**inp)
            e1.record()
            t3 = time.perf_counter()
            torch.cuda.synchronize()
            t4 = time.perf_counter()
            model.output_adapter(out)
            t5 = time.perf_counter()
            dev_ms += e0.elapsed_time(e1)
            for k, v in zip(("batch", "input_adapter", "forward_launch", "forward_wait", "output_adapter"), (t1 - t0, t2 - t1, t3 - t2, t4 - t3, t5 - t4)):
                acc[k] += v * 1e3
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(args.reps):
            model.run(**sample)
        torch.cuda.synchronize()
        acc["run_total"] = (time.perf_counter() - t0) * 1e3
    rec = {k + "_ms": v / args.reps for k, v in acc.items()}
    rec["forward_device_ms"] = dev_ms / args.reps
    rec["source_views"] = args.views
    print(json.dumps(rec))


if __name__ == "__main__":
    main()
