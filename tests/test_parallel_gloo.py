"""CPU, world size 2, ``gloo``: the depth-slab host logic of ``robustmvd_b200/parallel.py`` (partition, padded
all-gather along the plane axis, softargmax partial -> all-gather -> merge -> confidence all-reduce).

No GPU here, so the LOCAL compute of each rank is the oracle (injected through the ``ops`` / ``variance_fn`` /
block hooks that the module exposes for exactly this purpose); the collectives and the assembly are the product code.
The same functions run with the CUDA kernels as local compute in tests/test_gpu_parallel.py."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle
from robustmvd_b200 import parallel
from tools import cases

WORLD = 2


def test_slab_range_is_a_balanced_partition():
    for S in (1, 7, 64, 255, 256):
        for G in (1, 2, 3, 8):
            cuts = [parallel.slab_range(S, G, g) for g in range(G)]
            assert cuts[0][0] == 0 and cuts[-1][1] == S
            assert all(a[1] == b[0] for a, b in zip(cuts[:-1], cuts[1:]))
            lens = [hi - lo for lo, hi in cuts]
            assert max(lens) - min(lens) <= 1 and lens == sorted(lens, reverse=True)


class OracleBlock:
    """Stands in for PlanesweepCorrelation on CPU: same forward contract incl. ``plane_range``."""
    warp_only = False

    def get_sampling_invdepths(self, num_sampling_points, min_depth, max_depth, sampling_invdepths, sampling_type,
                               device=None):
        inv = oracle.compute_sampling_invdepths(min_depth, max_depth, num_sampling_points, sampling_type)
        return torch.from_numpy(inv)[:, :, None, None]

    def __call__(self, plane_range=None, **kw):
        npy = lambda t: t.numpy() if isinstance(t, torch.Tensor) else t  # noqa: E731
        corrs, masks, inv = oracle.planesweep_correlation(
            npy(kw["feat_key"]), npy(kw["intrinsics_key"]), [npy(f) for f in kw["feat_sources"]],
            [npy(t) for t in kw["source_to_key_transforms"]], None, num_sampling_points=kw["num_sampling_points"],
            min_depth=kw["min_depth"], max_depth=kw["max_depth"])
        lo, hi = plane_range
        return ([torch.from_numpy(c[:, lo:hi].copy()) for c in corrs], [torch.from_numpy(m[:, lo:hi].copy()) for m in masks],
                torch.from_numpy(inv)[:, :, None, None])


class OracleOps:
    @staticmethod
    def partial(cost_slab, depth_slab, d_begin):
        return torch.from_numpy(np.stack(oracle.softargmax_partials(cost_slab.numpy(), depth_slab.numpy(), d_begin)))

    @staticmethod
    def combine(parts):
        p = parts.numpy()
        return torch.from_numpy(np.stack(oracle.softargmax_combine([tuple(p[g]) for g in range(p.shape[0])])))

    @staticmethod
    def confidence(cost_slab, merged, D, d_begin):
        m = merged.numpy()
        return torch.from_numpy(oracle.softargmax_confidence_part(cost_slab.numpy(), m[0], m[1], m[3], D, d_begin))


def _variance_fn(feats, proj, depth, plane_range):
    lo, hi = plane_range
    return torch.from_numpy(oracle.mvs_variance(feats.numpy(), proj.numpy(), depth.numpy()[:, lo:hi]))


def _worker(rank, port, results):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=WORLD)
    try:
        t = torch.from_numpy
        # 1. plane-sweep correlation, S = 7 planes over 2 ranks (uneven slabs 4 + 3)
        case = [c for c in cases.planesweep_cases(False) if c["name"] == "odd_sizes"][0]
        kw = dict(feat_key=t(case["feat_key"]), intrinsics_key=t(case["intrinsics_key"]),
                  feat_sources=[t(f) for f in case["feat_sources"]],
                  source_to_key_transforms=[t(x) for x in case["transforms"]], **case["kwargs"])
        corrs, masks, inv = parallel.planesweep_correlation_slabs(OracleBlock(), **kw)
        oc, om, oinv = oracle.planesweep_correlation(case["feat_key"], case["intrinsics_key"], case["feat_sources"],
                                                     case["transforms"], None, **case["kwargs"])
        ok1 = all(np.array_equal(c.numpy(), o) for c, o in zip(corrs, oc)) and \
            all(np.array_equal(m.numpy(), o) for m, o in zip(masks, om)) and np.array_equal(inv.numpy()[:, :, 0, 0], oinv)
        slab_c, _, _ = parallel.planesweep_correlation_slabs(OracleBlock(), gather=False, **kw)
        lo, hi = parallel.slab_range(7, WORLD, rank)
        ok1 = ok1 and np.array_equal(slab_c[0].numpy(), oc[0][:, lo:hi])
        # 2. MVSNet variance: sharded and gathered
        m = cases.mvs_cases(False)[2]   # mvs_odd: D = 5 -> slabs 3 + 2
        full = oracle.mvs_variance(m["feats"], m["proj"], m["depth"])
        slab, (lo, hi) = parallel.mvs_variance_slabs(t(m["feats"]), t(m["proj"]), t(m["depth"]), variance_fn=_variance_fn)
        gathered, _ = parallel.mvs_variance_slabs(t(m["feats"]), t(m["proj"]), t(m["depth"]), gather=True,
                                                  variance_fn=_variance_fn)
        ok2 = np.array_equal(slab.numpy(), full[:, :, lo:hi]) and np.array_equal(gathered.numpy(), full)
        # 3. regression with D partitioned
        m = cases.mvs_cases(False)[0]
        D = m["cost"].shape[1]
        lo, hi = parallel.slab_range(D, WORLD, rank)
        r = parallel.mvs_regress_slabs(t(m["cost"][:, lo:hi].copy()), t(m["depth"][:, lo:hi].copy()), lo, D, ops=OracleOps)
        od, ocf, oi, oe = oracle.mvs_regress(m["cost"], m["depth"])
        near = np.abs(oe - np.round(oe)) < 1e-3
        same = r["depth_index"].numpy() == oi
        ok3 = bool(np.allclose(r["depth"].numpy(), od, rtol=1e-5)) and bool((same | near).all()) and \
            bool(np.abs(r["confidence"].numpy() - ocf)[same].max() < 1e-5) and bool(same.mean() > 0.9)
        results[rank] = (ok1, ok2, ok3)
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.timeout(300)
def test_depth_slab_path_world_size_2():
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        results = mgr.dict()
        port = _free_port()
        procs = [ctx.Process(target=_worker, args=(r, port, results)) for r in range(WORLD)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(240)
        assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
        assert dict(results) == {0: (True, True, True), 1: (True, True, True)}
