"""configs[0] of BASELINE.json: the UNMODIFIED reference ``robust_mvd`` on the reference's own ``sample_data`` (key view + 6 source
views, 720x1280), CPU, ``num_gpus=0`` (route B, tools/ref_models.py), frozen as a golden fixture.

    python -m tools.make_sample_data_golden          # container only (/root/reference)

``inference.load_data`` (reference ``inference.py:18-55``) reads the sample; the model is built with ``torch.manual_seed`` + its
own ``init_weights`` (the pretrained weights are URLs, unreachable here).  Written to tests/golden/sample_data.npz:
  png_v (file bytes of the 7 images), intrinsics [7,3,3], poses [7,4,4]  the inputs, exactly as load_data returns them
  pred_depth, aux_invdepth                                            model.run outputs
  mask_sha_v, mask_count_v, mask_rowsum_v / corr_sub_v, corr_absmax_v   corr_block.masks (sha256 of the packed bits, count, per-(plane,row)
                                                                      sums) / .corrs (stride-251 subsample + max |corr|) of the 6 views
  cpu_ms_model_run, cpu_ms_corr_block, cpu_threads                    wall times of the reference on this container's cores
"""
import hashlib
import importlib.util
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tools.ref_models import import_rmvd, REF_ROOT  # noqa: E402
from tools import cases  # noqa: E402

SUB = 251                  # stride of the corr subsample (prime)
OUT = os.path.join(ROOT, "tests", "golden", "sample_data.npz")


def main():
    rmvd = import_rmvd()
    spec = importlib.util.spec_from_file_location("_ref_inference", os.path.join(REF_ROOT, "inference.py"))
    inference = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(inference)
    sample, h_orig, w_orig = inference.load_data(os.path.join(REF_ROOT, "sample_data"))
    torch.set_num_threads(os.cpu_count())
    torch.manual_seed(cases.MODEL_SEED)
    model = rmvd.create_model("robust_mvd", pretrained=False, weights=None, train=False, num_gpus=0)
    pred, aux = model.run(**sample)          # warm-up (first call pays allocations)
    ms_corr = [0.0]
    inner = model.corr_block.forward

    def timed_forward(*a, **k):
        t = time.perf_counter()
        r = inner(*a, **k)
        ms_corr[0] = (time.perf_counter() - t) * 1e3
        return r
    model.corr_block.forward = timed_forward
    t0 = time.perf_counter()
    pred, aux = model.run(**sample)
    ms_run = (time.perf_counter() - t0) * 1e3
    corrs = [c.detach().numpy() for c in model.corr_block.corrs]
    masks = [m.detach().numpy() for m in model.corr_block.masks]
    root = os.path.join(REF_ROOT, "sample_data")
    pngs = [os.path.join(root, "key", "image.png")] + [os.path.join(root, "source", d, "image.png") for d in sorted(os.listdir(os.path.join(root, "source")))]
    out = dict(intrinsics=np.stack(sample["intrinsics"]).astype(np.float32), poses=np.stack(sample["poses"]).astype(np.float32),
               pred_depth=np.asarray(pred["depth"]), aux_invdepth=np.asarray(aux["invdepth"]), n_views=np.int32(len(corrs)),
               feat_shape=np.asarray(corrs[0].shape, np.int32), cpu_ms_model_run=np.float32(ms_run), cpu_threads=np.int32(torch.get_num_threads()))
    for v, path in enumerate(pngs):                      # the images as the PNG files themselves (smaller than any re-encoding)
        out[f"png_{v}"] = np.frombuffer(open(path, "rb").read(), dtype=np.uint8)
    for v, (c, m) in enumerate(zip(corrs, masks)):
        out[f"mask_sha_{v}"] = np.frombuffer(hashlib.sha256(np.packbits(m.astype(np.uint8)).tobytes()).digest(), dtype=np.uint8)
        out[f"mask_count_{v}"] = np.int64(m.sum())
        out[f"mask_rowsum_{v}"] = m.sum(axis=(0, 3)).astype(np.int32)          # [S, h]: localises a mismatch
        out[f"corr_sub_{v}"] = c.reshape(-1)[::SUB].copy()
        out[f"corr_absmax_{v}"] = np.float32(np.abs(c).max())
    out["cpu_ms_corr_block"] = np.float32(ms_corr[0])
    np.savez_compressed(OUT, **out)
    print({k: (getattr(v, "shape", None), str(getattr(v, "dtype", type(v)))) for k, v in out.items() if not k.startswith(("mask_", "corr_", "png_"))})
    print("model.run: %.0f ms (corr_block.forward %.0f ms) on %d threads; file %.1f MB" % (ms_run, ms_corr[0], torch.get_num_threads(), os.path.getsize(OUT) / 1e6))


if __name__ == "__main__":
    main()
