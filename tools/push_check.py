"""Depth-slab "push" transport against the single-GPU volumes and the other transports (cfg4, 1 frame).
    torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/push_check.py
Prints one JSON line per (views, variant) on rank 0."""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from robustmvd_b200 import parallel, synthetic as syn  # noqa: E402
from robustmvd_b200.blocks.planesweep_corr import PlanesweepCorrelation  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0)))
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    cfg = syn.CFG4
    cu = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)  # noqa: E731

    def timed(fn, steps=20, warmup=3):
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
        ms = torch.tensor([e0.elapsed_time(e1) / steps], dtype=torch.float64, device=dev)
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    for V in (2, 10):
        inp = syn.planesweep_inputs(cfg["h"], cfg["w"], cfg["C"], V, seed=77)
        kw = dict(feat_key=cu(inp["feat_key"]), intrinsics_key=cu(inp["intrinsics_key"]), feat_sources=[cu(f) for f in inp["feat_sources"]],
                  source_to_key_transforms=[cu(t) for t in inp["transforms"]], num_sampling_points=cfg["S"], min_depth=cfg["min_depth"],
                  max_depth=cfg["max_depth"])
        blk = PlanesweepCorrelation()
        full_c, full_m, _ = blk(**kw)
        ms_full = timed(lambda: blk(**kw))
        ms_slab = timed(lambda: blk(plane_range=parallel.slab_range(cfg["S"], world, rank), **kw))
        for name, args in (("stores", dict(transport="stores")), ("push4", dict(transport="push", push_ctas=4)),
                           ("push8", dict(transport="push", push_ctas=8)), ("push16", dict(transport="push", push_ctas=16)),
                           ("push32", dict(transport="push", push_ctas=32))):
            gather = parallel.SlabGather(blk, V, 1, cfg["S"], cfg["h"], cfg["w"], dev, **args)
            oks = []
            for _ in range(3):                       # both buffer sets and a reused queue
                corrs, masks, _ = gather(**kw)
                torch.cuda.synchronize()
                oks.append(all(torch.equal(a, b) for a, b in zip(corrs, full_c)) and all(torch.equal(a, b) for a, b in zip(masks, full_m)))
            okt = torch.tensor([int(all(oks))], device=dev)
            dist.all_reduce(okt, op=dist.ReduceOp.MIN)
            ms = timed(lambda: gather(**kw))
            if rank == 0:
                print(json.dumps({"views": V, "n_gpus": world, "variant": name, "mode": gather.mode, "bit_exact": bool(okt.item()),
                                  "ms_total": ms, "ms_1gpu": ms_full, "ms_slab_compute_only": ms_slab, "speedup": ms_full / ms}), flush=True)
            del gather
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
