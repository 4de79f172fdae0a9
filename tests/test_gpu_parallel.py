"""GPU: ``robustmvd_b200.parallel`` with the CUDA kernels as the local compute.  On one GPU the process group has a
single rank (NCCL, 127.0.0.1); the multi-rank assembly itself is covered on CPU by tests/test_parallel_gloo.py and on
2+ GPUs by ``torchrun tools/slab_check.py`` (run through ``gpurun --gpus N``)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist

import oracle
from robustmvd_b200 import parallel, synthetic as syn
from tools import cases

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def nccl_single_rank():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(0)
    dist.init_process_group("nccl", rank=0, world_size=1, device_id=torch.device("cuda", 0))
    yield
    dist.destroy_process_group()


def test_slab_functions_single_rank(cuda_lib, nccl_single_rank):
    from robustmvd_b200.blocks.planesweep_corr import PlanesweepCorrelation
    from robustmvd_b200.blocks import utils as bu
    cu = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
    inp = syn.planesweep_inputs(24, 40, 64, 2, seed=3)
    kw = dict(feat_key=cu(inp["feat_key"]), intrinsics_key=cu(inp["intrinsics_key"]),
              feat_sources=[cu(f) for f in inp["feat_sources"]], source_to_key_transforms=[cu(t) for t in inp["transforms"]],
              num_sampling_points=37, min_depth=0.4, max_depth=1000.0)
    corrs, masks, inv = parallel.planesweep_correlation_slabs(PlanesweepCorrelation(), **kw)
    c0, m0, _ = PlanesweepCorrelation()(**kw)
    assert all(torch.equal(a, b) for a, b in zip(corrs, c0)) and all(torch.equal(a, b) for a, b in zip(masks, m0))
    m = cases.mvs_cases(False)[0]
    r = parallel.mvs_regress_slabs(cu(m["cost"]), cu(m["depth"]), 0, m["cost"].shape[1])
    od, oc, oi, oe = oracle.mvs_regress(m["cost"], m["depth"])
    assert (np.abs(r["depth"].cpu().numpy() - od) <= 1e-3 * np.abs(od)).all()
    near = np.abs(oe - np.round(oe)) < 1e-3
    same = r["depth_index"].cpu().numpy() == oi
    assert (same | near).all() and np.abs(r["confidence"].cpu().numpy() - oc)[same].max() < 1e-4
    slab, (lo, hi) = parallel.mvs_variance_slabs(cu(m["feats"]), cu(m["proj"]), cu(m["depth"]))
    assert (lo, hi) == (0, m["depth"].shape[1])
    assert torch.equal(slab, bu.mvs_variance_volume(cu(m["feats"]), cu(m["proj"]), cu(m["depth"])))


def test_emulated_ranks_assemble_to_the_full_volume(cuda_lib):
    """G ranks emulated one after another on one GPU: their slabs, concatenated in rank order, are the full volume."""
    from robustmvd_b200.blocks.planesweep_corr import PlanesweepCorrelation
    cu = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
    inp = syn.planesweep_inputs(24, 40, 64, 2, seed=5)
    kw = dict(feat_key=cu(inp["feat_key"]), intrinsics_key=cu(inp["intrinsics_key"]),
              feat_sources=[cu(f) for f in inp["feat_sources"]], source_to_key_transforms=[cu(t) for t in inp["transforms"]],
              num_sampling_points=61, min_depth=0.4, max_depth=1000.0)
    blk = PlanesweepCorrelation()
    full_c, full_m, _ = blk(**kw)
    full_c, full_m = [c.clone() for c in full_c], [m.clone() for m in full_m]
    for G in (2, 3, 8):
        parts = [blk(plane_range=parallel.slab_range(61, G, g), **kw)[:2] for g in range(G)]
        parts = [([c.clone() for c in p[0]], [m.clone() for m in p[1]]) for p in parts]
        for v in range(2):
            assert torch.equal(torch.cat([p[0][v] for p in parts], 1), full_c[v])
            assert torch.equal(torch.cat([p[1][v] for p in parts], 1), full_m[v])


@pytest.mark.parametrize("variant,w", [("planes", 40), ("tiles", 40), ("push", 40), ("push", 42)])
@pytest.mark.parametrize("G", [2, 3])
def test_peer_transports_with_emulated_ranks(cuda_lib, variant, w, G):
    """The depth-slab kernels of ``parallel.SlabGather`` (rmvd_psweep_corr_fwd_peers with the plane and the tile partition,
    rmvd_psweep_corr_fwd_push) with G ranks emulated one after another on one GPU: every "rank" owns a [2,V,N,S,h,w] buffer,
    all of them addressable by all (the same device), and the flags of the closing handshake are preset so that no rank waits
    for one that has not run yet.  After the G calls every buffer holds the full volumes, bit for bit those of one plain call."""
    from robustmvd_b200.blocks.planesweep_corr import PlanesweepCorrelation
    cu = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
    h, V, S = 24, 3, 37                                  # (w = 42: rows that are not 16-byte multiples, the pusher's scalar path)
    inp = syn.planesweep_inputs(h, w, 64, V, seed=11)
    kw = dict(feat_key=cu(inp["feat_key"]), intrinsics_key=cu(inp["intrinsics_key"]),
              feat_sources=[cu(f) for f in inp["feat_sources"]], source_to_key_transforms=[cu(t) for t in inp["transforms"]],
              num_sampling_points=S, min_depth=0.4, max_depth=1000.0)
    blk = PlanesweepCorrelation()
    full_c, full_m, _ = blk(**kw)
    N = full_c[0].shape[0]
    lib = cuda_lib
    side = torch.cuda.Stream()
    for epoch in (1, 2):                                 # the second call reuses flags, queues and counters
        bufs = [torch.full((2, V, N, S, h, w), float("nan"), device="cuda") for _ in range(G)]
        flags = [torch.full((64,), epoch, dtype=torch.int32, device="cuda") for _ in range(G)]
        if epoch == 1:
            queues = []
            for r in range(G):
                lo, hi = parallel.slab_range(S, G, r)
                nb = lib.rmvd_psweep_push_queue_bytes(N, h, w, V, max(hi - lo, 1))
                assert nb > 16
                queues.append(torch.zeros((nb + 15) // 16 * 4, dtype=torch.int32, device="cuda"))
        half = bufs[0][0].numel() * 4
        for r in range(G):
            peers = dict(rank=r, epoch=epoch, corr=[b.data_ptr() for b in bufs], mask=[b.data_ptr() + half for b in bufs],
                         flags=[f.data_ptr() for f in flags])
            lo, hi = parallel.slab_range(S, G, r)
            if variant == "tiles":
                lo, hi = 0, S
                peers["partition"] = 1
            if variant == "push":
                peers.update(queue=queues[r], push_stream=side, push_ctas=4)
            blk._forward_impl(plane_range=(lo, hi), peers=peers, **kw)
            torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        for b in bufs:
            assert all(torch.equal(b[0, v], full_c[v]) for v in range(V)) and all(torch.equal(b[1, v], full_m[v]) for v in range(V))
        assert all(int(f[r].item()) == epoch for f in flags for r in range(G))
