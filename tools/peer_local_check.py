"""One GPU: rmvd_psweep_corr_fwd_peers with both 'peers' pointing at LOCAL buffers (flags pre-set so that the handshake passes),
to separate the cost of the peer-store code path from the cost of the NVLink stores themselves.
    python tools/peer_local_check.py [--views V]"""
import argparse
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from robustmvd_b200 import parallel, synthetic as syn  # noqa: E402
from robustmvd_b200.blocks.planesweep_corr import PlanesweepCorrelation  # noqa: E402


def timed(fn, steps=30, warmup=5):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--views", type=int, default=2)
    ap.add_argument("--peers", type=int, default=2)
    a = ap.parse_args()
    cfg, V, G = syn.CFG4, a.views, a.peers
    dev = "cuda:0"
    cu = lambda x: torch.from_numpy(np.ascontiguousarray(x)).to(dev)  # noqa: E731
    inp = syn.planesweep_inputs(cfg["h"], cfg["w"], cfg["C"], V, seed=77)
    kw = dict(feat_key=cu(inp["feat_key"]), intrinsics_key=cu(inp["intrinsics_key"]), feat_sources=[cu(f) for f in inp["feat_sources"]],
              source_to_key_transforms=[cu(t) for t in inp["transforms"]], num_sampling_points=cfg["S"], min_depth=cfg["min_depth"],
              max_depth=cfg["max_depth"])
    blk = PlanesweepCorrelation()
    S = cfg["S"]
    bufs = [torch.zeros((2, V, 1, S, cfg["h"], cfg["w"]), device=dev) for _ in range(G)]
    flags = [torch.full((64,), 0x7fffff00, dtype=torch.int32, device=dev) for _ in range(G)]
    half = bufs[0][0].numel() * 4
    out = {"views": V, "fake_peers": G}
    out["ms_full"] = timed(lambda: blk(**kw))
    lo, hi = parallel.slab_range(S, G, 0)
    out["ms_plane_slab_plain"] = timed(lambda: blk(plane_range=(lo, hi), **kw))
    epoch = [0]

    def call(partition):
        epoch[0] += 1
        peers = dict(rank=0, epoch=epoch[0], corr=[b.data_ptr() for b in bufs], mask=[b.data_ptr() + half for b in bufs],
                     flags=[f.data_ptr() for f in flags], partition=partition)
        pr = (lo, hi) if partition == 0 else (0, S)
        blk._forward_impl(plane_range=pr, peers=peers, **kw)

    out["ms_plane_slab_local_peers"] = timed(lambda: call(0))
    out["ms_tile_part_local_peers"] = timed(lambda: call(1))
    print(json.dumps(out))


if __name__ == "__main__":
    main()
