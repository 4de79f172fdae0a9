"""Multi-GPU check + timing of the depth-slab partition (BASELINE.json configs[3]: robust_mvd at the ETH3D shape
768x1152 -> 96x144 features, C=256, S=256, slabs over 2/4/8 B200).  Launch on the GPU box:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 \
        tools/slab_check.py [--views 2] [--steps 20]

Every rank holds the full features, computes its slab of planes and all-gathers corr/mask (NCCL over NVLink).  Rank 0
verifies the gathered volumes against a full single-GPU call bit for bit and prints one JSON line with the device time
(max over ranks) of slab compute alone and of compute + gather, next to the single-GPU time.
"""
import argparse
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from robustmvd_b200 import parallel, synthetic as syn  # noqa: E402
from robustmvd_b200.blocks.planesweep_corr import PlanesweepCorrelation  # noqa: E402
from robustmvd_b200.blocks import utils as bu  # noqa: E402


def timed(fn, steps, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / steps], dtype=torch.float64, device="cuda")
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--views", type=int, default=2)
    ap.add_argument("--steps", type=int, default=20)
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    cfg = syn.CFG4
    inp = syn.planesweep_inputs(cfg["h"], cfg["w"], cfg["C"], args.views, seed=77)   # same seed: replicated features
    cu = lambda a: torch.from_numpy(a).cuda()  # noqa: E731
    kw = dict(feat_key=cu(inp["feat_key"]), intrinsics_key=cu(inp["intrinsics_key"]),
              feat_sources=[cu(f) for f in inp["feat_sources"]], source_to_key_transforms=[cu(t) for t in inp["transforms"]],
              num_sampling_points=cfg["S"], min_depth=cfg["min_depth"], max_depth=cfg["max_depth"])
    blk = PlanesweepCorrelation()
    corrs, masks, _ = parallel.planesweep_correlation_slabs(blk, **kw)
    full_c, full_m, _ = PlanesweepCorrelation()(**kw)
    ok = all(torch.equal(a, b) for a, b in zip(corrs, full_c)) and all(torch.equal(a, b) for a, b in zip(masks, full_m))
    # regression with D partitioned (cfg3 cost shape)
    m = syn.mvs_inputs(syn.CFG3["h"], syn.CFG3["w"], 4, 2, syn.CFG3["D"], seed=5)
    D = syn.CFG3["D"]
    lo, hi = parallel.slab_range(D, world, rank)
    cost, depth = cu(m["cost"]), cu(m["depth"])
    cs, ds = cost[:, lo:hi].contiguous(), depth[:, lo:hi].contiguous()
    r = parallel.mvs_regress_slabs(cs, ds, lo, D)
    r0 = bu.mvs_regress(cost, depth)
    rel = ((r["depth"] - r0["depth"]).abs() / r0["depth"].abs()).max().item()
    idx_same = (r["depth_index"] == r0["depth_index"]).float().mean().item()
    conf_err = (r["confidence"] - r0["confidence"]).abs()[r["depth_index"] == r0["depth_index"]].max().item()
    ok_all = torch.tensor([int(ok and rel < 1e-3 and idx_same > 0.999 and conf_err < 1e-4)], device="cuda")
    dist.all_reduce(ok_all, op=dist.ReduceOp.MIN)

    t_full = timed(lambda: blk(**kw), args.steps)
    t_slab = timed(lambda: blk(plane_range=parallel.slab_range(cfg["S"], world, rank), **kw), args.steps)
    t_gather = timed(lambda: parallel.planesweep_correlation_slabs(blk, **kw), args.steps)
    t_reg_full = timed(lambda: bu.mvs_regress(cost, depth), args.steps)
    t_reg_slab = timed(lambda: parallel.mvs_regress_slabs(cs, ds, lo, D), args.steps)
    if rank == 0:
        vox = args.views * cfg["S"] * cfg["h"] * cfg["w"]
        print(json.dumps({"check": "depth_slabs", "n_gpus": world, "views": args.views, "parity_ok": bool(ok_all.item()),
                          "regress_rel_err": rel, "regress_index_agreement": idx_same, "regress_conf_err": conf_err,
                          "ms_full_1gpu": t_full, "ms_slab_compute": t_slab, "ms_slab_compute_plus_allgather": t_gather,
                          "voxels_per_s_slabs": vox / (t_gather * 1e-3), "voxels_per_s_1gpu": vox / (t_full * 1e-3),
                          "ms_regress_full_1gpu": t_reg_full, "ms_regress_slabs": t_reg_slab}), flush=True)
    dist.destroy_process_group()
    sys.exit(0 if ok_all.item() else 1)


if __name__ == "__main__":
    main()
