"""GPU parity: the CUDA plane-sweep correlation (through the C ABI) against the numpy oracle and the golden vectors
frozen from the unmodified reference.  Bars: masks, geometric masks and integer sampling indices bit-exact; corr within
1e-4 of max|ref| (BASELINE.json north_star)."""
import numpy as np
import pytest
import torch

import oracle
from tools import cases
from robustmvd_b200 import synthetic as syn
from golden_util import load_golden, check_planesweep_against_golden

pytestmark = pytest.mark.gpu
PS = cases.planesweep_cases(include_large=True)
TOL = 1e-4  # relative to max|corr| of the view


def _run_cuda(case, plane_range=None, debug=True):
    from robustmvd_b200.blocks.planesweep_corr import PlanesweepCorrelation
    dev = "cuda:0"
    t = lambda a: None if a is None else torch.from_numpy(np.asarray(a)).to(dev)  # noqa: E731
    kw = {k: (t(v) if isinstance(v, np.ndarray) else v) for k, v in case["kwargs"].items()}
    blk = PlanesweepCorrelation(warp_only=case.get("warp_only", False), normalize=case.get("normalize", "dim"))
    corrs, masks, inv = blk(
        feat_key=t(case["feat_key"]), intrinsics_key=t(case["intrinsics_key"]),
        feat_sources=[t(f) for f in case["feat_sources"]], source_to_key_transforms=[t(x) for x in case["transforms"]],
        intrinsics_sources=None if case["intrinsics_sources"] is None else [t(x) for x in case["intrinsics_sources"]],
        plane_range=plane_range, debug=debug and not case.get("warp_only", False), **kw)
    torch.cuda.synchronize()
    dbg = blk.debug
    return ([c.cpu().numpy() for c in corrs], [m.cpu().numpy() for m in masks], inv.cpu().numpy(),
            None if dbg is None else {k: v.cpu().numpy() for k, v in dbg.items()})


@pytest.mark.parametrize("case", PS, ids=[c["name"] for c in PS])
def test_cuda_matches_oracle_and_golden(cuda_lib, case):
    corrs, masks, inv, dbg = _run_cuda(case)
    ocorrs, omasks, oinv, odbg = oracle.planesweep_correlation(
        case["feat_key"], case["intrinsics_key"], case["feat_sources"], case["transforms"], case["intrinsics_sources"],
        warp_only=case.get("warp_only", False), normalize=case.get("normalize", "dim"), return_debug=True,
        **case["kwargs"])
    assert np.array_equal(inv, oinv)
    for v in range(len(corrs)):
        assert np.array_equal(masks[v], omasks[v]), f"mask view {v}: {(masks[v] != omasks[v]).sum()} mismatches"
        scale = max(float(np.abs(ocorrs[v]).max()), 1e-30)
        err = float(np.abs(corrs[v] - ocorrs[v]).max()) / scale
        assert err <= TOL, f"corr view {v}: {err}"
        if dbg is not None:
            assert np.array_equal(dbg["ix0"][v], odbg[v]["ix0"])
            assert np.array_equal(dbg["iy0"][v], odbg[v]["iy0"])
            assert np.array_equal(dbg["geom"][v].astype(bool), odbg[v]["geom_mask"])
    g = load_golden("ps", case["name"])
    if dbg is not None:
        check_planesweep_against_golden(g, corrs, masks, inv, list(dbg["ix0"]), list(dbg["iy0"]), list(dbg["geom"]), TOL)
    else:
        check_planesweep_against_golden(g, corrs, masks, inv, tol=TOL)


# key map (h, w), source map (hs, ws), channels: every kind of tile rectangle the launcher picks (row segments, 32 x 2, whole
# rows of narrow maps, one-pixel-wide maps, ragged last tiles) and source maps of another size
TILE_SHAPES = [(7, 33, 7, 33, 32), (10, 130, 10, 130, 64), (100, 8, 100, 8, 32), (50, 3, 50, 3, 32), (70, 1, 70, 1, 32),
               (2, 200, 2, 200, 32), (1, 96, 1, 96, 64), (31, 31, 17, 45, 32), (9, 65, 20, 130, 32), (5, 64, 5, 64, 96)]


@pytest.mark.parametrize("shape", TILE_SHAPES, ids=["%dx%d_src%dx%d_c%d" % s for s in TILE_SHAPES])
def test_tile_shapes_against_oracle(cuda_lib, shape):
    """Odd map sizes through the tensor-core kernel: masks and integer indices bit-exact, corr within tolerance."""
    h, w, hs, ws, C = shape
    inp = syn.planesweep_inputs(h, w, C, 2, n=2, seed=500 + h + w, hs=hs, ws=ws)
    case = dict(name="shape", feat_key=inp["feat_key"], intrinsics_key=inp["intrinsics_key"],
                feat_sources=inp["feat_sources"], transforms=inp["transforms"], intrinsics_sources=None,
                kwargs=dict(num_sampling_points=40, min_depth=0.4, max_depth=1000.0))
    corrs, masks, inv, dbg = _run_cuda(case)
    ocorrs, omasks, oinv, odbg = oracle.planesweep_correlation(
        case["feat_key"], case["intrinsics_key"], case["feat_sources"], case["transforms"], None, return_debug=True,
        **case["kwargs"])
    for v in range(2):
        assert np.array_equal(masks[v], omasks[v]), f"mask view {v}: {(masks[v] != omasks[v]).sum()} mismatches"
        assert np.array_equal(dbg["ix0"][v], odbg[v]["ix0"]) and np.array_equal(dbg["iy0"][v], odbg[v]["iy0"])
        assert np.array_equal(dbg["geom"][v].astype(bool), odbg[v]["geom_mask"])
        scale = max(float(np.abs(ocorrs[v]).max()), 1e-30)
        assert float(np.abs(corrs[v] - ocorrs[v]).max()) / scale <= TOL
    # same call without the debug outputs (dead-plane skipping enabled)
    corrs2, masks2, _, _ = _run_cuda(case, debug=False)
    for v in range(2):
        assert np.array_equal(masks2[v], masks[v]) and np.array_equal(corrs2[v], corrs[v])


def test_randomised_cases_against_oracle(cuda_lib, monkeypatch):
    """tools/tc_fuzz.py: random shapes, poses, plane counts up to 300 and explicit schedules with zero / negative / inf / nan
    entries; masks and indices bit-exact, with and without the debug outputs."""
    from tools import tc_fuzz
    monkeypatch.setattr("sys.argv", ["tc_fuzz", "40", "11"])
    assert tc_fuzz.main() == 0


def test_depth_slabs_concatenate_to_full_volume(cuda_lib):
    case = [c for c in PS if c["name"] == "kitti_like_c32"][0]
    full_c, full_m, _, _ = _run_cuda(case, debug=False)
    S = case["kwargs"]["num_sampling_points"]
    cuts = [0, 37, 128, 129, S]
    for v in range(len(full_c)):
        parts_c, parts_m = [], []
        for lo, hi in zip(cuts[:-1], cuts[1:]):
            c, m, _, _ = _run_cuda(case, plane_range=(lo, hi), debug=False)
            parts_c.append(c[v])
            parts_m.append(m[v])
        assert np.array_equal(np.concatenate(parts_c, 1), full_c[v])
        assert np.array_equal(np.concatenate(parts_m, 1), full_m[v])


def test_cfg2_full_size_against_oracle(cuda_lib):
    """BASELINE.json config 2 at full size: KITTI 48x160 features, C=256, S=256, 2 source views."""
    cfg = syn.CFG2
    inp = syn.planesweep_inputs(cfg["h"], cfg["w"], cfg["C"], cfg["V"])
    case = dict(name="cfg2", feat_key=inp["feat_key"], intrinsics_key=inp["intrinsics_key"],
                feat_sources=inp["feat_sources"], transforms=inp["transforms"], intrinsics_sources=None,
                kwargs=dict(num_sampling_points=cfg["S"], min_depth=cfg["min_depth"], max_depth=cfg["max_depth"]))
    corrs, masks, inv, dbg = _run_cuda(case)
    ocorrs, omasks, oinv, odbg = oracle.planesweep_correlation(
        case["feat_key"], case["intrinsics_key"], case["feat_sources"], case["transforms"], None, return_debug=True,
        **case["kwargs"])
    for v in range(2):
        assert np.array_equal(masks[v], omasks[v])
        assert np.array_equal(dbg["ix0"][v], odbg[v]["ix0"]) and np.array_equal(dbg["iy0"][v], odbg[v]["iy0"])
        scale = float(np.abs(ocorrs[v]).max())
        assert float(np.abs(corrs[v] - ocorrs[v]).max()) <= TOL * scale
        assert 0.3 < masks[v].mean() < 0.95  # the synthetic geometry keeps a useful share of samples in view


def test_linearity_and_mask_invariance_at_eth3d_size(cuda_lib):
    """Size-independent properties at BASELINE config 4 (96x144, C=256, S=256): corr is linear in the key features and
    the mask does not depend on the features at all."""
    cfg = syn.CFG4
    inp = syn.planesweep_inputs(cfg["h"], cfg["w"], cfg["C"], 2, seed=77)
    base = dict(name="cfg4", intrinsics_key=inp["intrinsics_key"], feat_sources=inp["feat_sources"],
                transforms=inp["transforms"], intrinsics_sources=None,
                kwargs=dict(num_sampling_points=cfg["S"], min_depth=cfg["min_depth"], max_depth=cfg["max_depth"]))
    k1 = inp["feat_key"]
    k2 = syn.features(k1.shape, 991)
    c1, m1, _, _ = _run_cuda(dict(base, feat_key=k1), debug=False)
    c2, m2, _, _ = _run_cuda(dict(base, feat_key=k2), debug=False)
    c3, m3, _, _ = _run_cuda(dict(base, feat_key=(2.0 * k1 - 0.5 * k2).astype(np.float32)), debug=False)
    for v in range(2):
        assert np.array_equal(m1[v], m2[v]) and np.array_equal(m1[v], m3[v])
        ref = 2.0 * c1[v] - 0.5 * c2[v]
        assert np.abs(c3[v] - ref).max() <= 1e-4 * np.abs(ref).max()
        assert (c1[v][m1[v] == 0] == 0).all()


def test_torchcorr_on_explicit_grids(cuda_lib):
    from robustmvd_b200.blocks.planesweep_corr import TorchCorr
    case = [c for c in PS if c["name"] == "src_res_differs"][0]
    _, _, _, odbg = oracle.planesweep_correlation(
        case["feat_key"], case["intrinsics_key"], case["feat_sources"], case["transforms"], case["intrinsics_sources"],
        return_debug=True, **case["kwargs"])
    for nm in ("dim", False, True):
        for v in range(2):
            us, vs, gm = odbg[v]["us"], odbg[v]["vs"], odbg[v]["geom_mask"]
            oc, om = oracle.torch_corr(case["feat_key"], case["feat_sources"][v], us, vs, mask=gm, normalize=nm)
            grids = torch.from_numpy(np.stack([us, vs], 2)).cuda()
            c, m = TorchCorr(normalize=nm)(torch.from_numpy(case["feat_key"]).cuda(),
                                           torch.from_numpy(case["feat_sources"][v]).cuda(), grids,
                                           torch.from_numpy(gm).cuda())
            assert np.array_equal(m.cpu().numpy(), om)
            assert np.abs(c.cpu().numpy() - oc).max() <= TOL * np.abs(oc).max()


def test_warp_function_zeros_and_border(cuda_lib):
    from robustmvd_b200.blocks.planesweep_corr import warp
    rng = np.random.default_rng(5)
    x = rng.standard_normal((3, 5, 9, 14), dtype=np.float32)
    grid = np.stack([rng.uniform(-3, 17, (3, 6, 8)), rng.uniform(-3, 12, (3, 6, 8))], 1).astype(np.float32)
    b = oracle.bilinear_setup(grid[:, 0], grid[:, 1], 9, 14)
    from oracle.planesweep import grid_sample_bilinear_zeros
    ref = grid_sample_bilinear_zeros(x, b)
    out, mask = warp(torch.from_numpy(x).cuda(), grid=torch.from_numpy(grid).cuda(), padding_mode="zeros")
    assert np.abs(out.cpu().numpy() - ref).max() < 1e-5
    assert np.array_equal(mask.cpu().numpy()[:, 0], b["samp_mask"])
    ref_b = torch.nn.functional.grid_sample(
        torch.from_numpy(x), torch.from_numpy(np.stack([2 * grid[:, 0] / 14 - 1, 2 * grid[:, 1] / 9 - 1], -1)),
        padding_mode="border", align_corners=False).numpy()
    out_b, mask_b = warp(torch.from_numpy(x).cuda(), grid=torch.from_numpy(grid).cuda(), padding_mode="border")
    assert np.abs(out_b.cpu().numpy() - ref_b).max() < 1e-4
    assert (mask_b.cpu().numpy() == 1).all()


def test_shared_reciprocal_division_is_exact(cuda_lib):
    """The exact coordinate chain divides through a shared refined reciprocal (common.cuh: rcp_refined /
    div_shared); within the guarded operand range it must equal the correctly rounded quotient bit for bit."""
    from robustmvd_b200 import _lib
    rng = np.random.default_rng(123)
    n = 1 << 22
    mant = lambda: rng.uniform(1.0, 2.0, n).astype(np.float32) * rng.choice(np.array([-1.0, 1.0], np.float32), n)  # noqa: E731
    a = mant() * np.exp2(rng.integers(-58, 59, n)).astype(np.float32)
    b = mant() * np.exp2(rng.integers(-58, 59, n)).astype(np.float32)
    sizes = np.array([1, 3, 7, 9, 14, 20, 40, 48, 96, 144, 160, 224, 304, 1216, 32767], np.float32)
    b[: n // 2] = rng.choice(sizes, n // 2)                      # the divisions by the source map size
    a[: n // 4] = rng.uniform(-700, 700, n // 4).astype(np.float32)  # pixel-range numerators
    a[0:4] = [0.0, -0.0, 1e-30, 2.0 ** 60]
    ta, tb = torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()
    q1, q2 = torch.empty_like(ta), torch.empty_like(ta)
    _lib.check(cuda_lib.rmvd_debug_div_check(_lib.ptr(ta), _lib.ptr(tb), n, _lib.ptr(q1), _lib.ptr(q2), _lib.stream_ptr()), "div")
    torch.cuda.synchronize()
    assert torch.equal(q1.view(torch.int32)[1:], q2.view(torch.int32)[1:]) or \
        int((q1.view(torch.int32) != q2.view(torch.int32)).sum()) <= 1   # only -0 / +0 of a zero numerator may differ
    assert float((q1 - q2).abs().max()) == 0.0


@pytest.mark.parametrize("compact", ["bits", "u8", False])
def test_host_pipeline_equals_direct_call(cuda_lib, compact):
    """HostPlanesweepPipeline (chunked H2D / kernels / D2H on three streams) returns the block's outputs bit for bit, with the
    mask sent as bytes and expanded on the host (compact_mask=True) or as float32."""
    from robustmvd_b200 import synthetic as syn
    from robustmvd_b200.blocks.planesweep_corr import PlanesweepCorrelation
    from robustmvd_b200.pipeline import HostPlanesweepPipeline
    N, V, S = 5, 2, 48
    inp = syn.planesweep_inputs(24, 40, 64, V, n=N, seed=91)
    kw = dict(num_sampling_points=S, min_depth=0.4, max_depth=1000.0)
    dev = torch.device("cuda:0")
    Kk = torch.from_numpy(inp["intrinsics_key"]).to(dev)
    Ts = [torch.from_numpy(x).to(dev) for x in inp["transforms"]]
    hk = torch.from_numpy(inp["feat_key"]).pin_memory()
    hs = [torch.from_numpy(f).pin_memory() for f in inp["feat_sources"]]
    out = torch.zeros((2, V, N, S, 24, 40), dtype=torch.float32).pin_memory()
    pipe = HostPlanesweepPipeline(dev, PlanesweepCorrelation(), chunk_frames=2, compact_mask=bool(compact), mask_threads=3,
                                  mask_transport=compact or "bits")
    for _ in range(2):   # second call reuses the persistent device buffers
        out.fill_(7.0)
        pipe.run(hk, Kk, hs, Ts, out[0], out[1], **kw)
    corrs, masks, _ = PlanesweepCorrelation()(feat_key=hk.to(dev), intrinsics_key=Kk, feat_sources=[f.to(dev) for f in hs],
                                              source_to_key_transforms=Ts, **kw)
    for v in range(V):
        assert torch.equal(out[0, v], corrs[v].cpu()) and torch.equal(out[1, v], masks[v].cpu())


@pytest.mark.gpu
@pytest.mark.parametrize("compact", ["bits", "u8", False])
def test_host_pipeline_stream_of_batches(cuda_lib, compact):
    """run(wait=False) per batch + finish(): consecutive batches overlap (double-buffered device inputs) and every batch
    still equals the direct call on its own inputs."""
    from robustmvd_b200 import synthetic as syn
    from robustmvd_b200.blocks.planesweep_corr import PlanesweepCorrelation
    from robustmvd_b200.pipeline import HostPlanesweepPipeline
    N, V, S, B = 3, 2, 40, 5
    kw = dict(num_sampling_points=S, min_depth=0.4, max_depth=1000.0)
    dev = torch.device("cuda:0")
    pipe = HostPlanesweepPipeline(dev, PlanesweepCorrelation(), chunk_frames=[1, 2], compact_mask=bool(compact), mask_transport=compact or "bits")
    batches = []
    for b in range(B):
        inp = syn.planesweep_inputs(24, 40, 64, V, n=N, seed=200 + b)
        batches.append(dict(
            Kk=torch.from_numpy(inp["intrinsics_key"]).to(dev), Ts=[torch.from_numpy(x).to(dev) for x in inp["transforms"]],
            hk=torch.from_numpy(inp["feat_key"]).pin_memory(), hs=[torch.from_numpy(f).pin_memory() for f in inp["feat_sources"]],
            out=torch.full((2, V, N, S, 24, 40), 7.0, dtype=torch.float32).pin_memory()))
    for x in batches:
        pipe.run(x["hk"], x["Kk"], x["hs"], x["Ts"], x["out"][0], x["out"][1], wait=False, **kw)
    pipe.finish()
    blk = PlanesweepCorrelation()
    for x in batches:
        corrs, masks, _ = blk(feat_key=x["hk"].to(dev), intrinsics_key=x["Kk"], feat_sources=[f.to(dev) for f in x["hs"]],
                              source_to_key_transforms=x["Ts"], **kw)
        for v in range(V):
            assert torch.equal(x["out"][0, v], corrs[v].cpu()) and torch.equal(x["out"][1, v], masks[v].cpu())


@pytest.mark.parametrize("C,normalize,h,w", [(64, "dim", 24, 40), (256, "dim", 13, 37), (96, True, 10, 70), (48, "dim", 8, 16)])
def test_fused_leaky_relu_equals_activation_then_call(cuda_lib, C, normalize, h, w):
    """rmvd_psweep_corr_fwd_act (the encoder's last LeakyReLU fused into the operand emission) on pre-activation maps
    equals torch's leaky_relu followed by the ordinary call, bit for bit, and hands back the activated key map; C = 48 takes the
    documented fallback (shape outside the persistent kernel)."""
    from robustmvd_b200.blocks.planesweep_corr import PlanesweepCorrelation
    N, V, S = 2, 2, 40
    inp = syn.planesweep_inputs(h, w, C, V, n=N, seed=300 + C)
    dev = torch.device("cuda:0")
    g = torch.Generator(device="cpu").manual_seed(C)
    pre_key = (torch.from_numpy(inp["feat_key"]) * torch.where(torch.rand(N, C, h, w, generator=g) < 0.5, -3.0, 1.0)).to(dev)
    pre_src = [(torch.from_numpy(f) * torch.where(torch.rand(f.shape, generator=g) < 0.5, -3.0, 1.0)).to(dev) for f in inp["feat_sources"]]
    pre_key[0, 0, 0, :4] = torch.tensor([0.0, -0.0, float("nan"), -1e-40])
    Kk = torch.from_numpy(inp["intrinsics_key"]).to(dev)
    Ts = [torch.from_numpy(x).to(dev) for x in inp["transforms"]]
    kw = dict(intrinsics_key=Kk, source_to_key_transforms=Ts, num_sampling_points=S, min_depth=0.4, max_depth=1000.0)
    blk = PlanesweepCorrelation(normalize=normalize)
    key_out = torch.full_like(pre_key, 7.0)
    c1, m1, _ = blk(feat_key=pre_key, feat_sources=pre_src, fused_leaky_relu=0.2, key_activated_out=key_out, **kw)
    c1, m1 = [c.clone() for c in c1], [m.clone() for m in m1]
    lr = torch.nn.functional.leaky_relu
    c2, m2, _ = blk(feat_key=lr(pre_key, 0.2), feat_sources=[lr(f, 0.2) for f in pre_src], **kw)
    assert torch.equal(key_out.view(torch.int32), lr(pre_key, 0.2).view(torch.int32))
    for v in range(V):
        assert torch.equal(m1[v], m2[v])
        assert torch.equal(c1[v].view(torch.int32), c2[v].view(torch.int32))
    with pytest.raises(Exception):
        with torch.enable_grad():
            blk(feat_key=pre_key.clone().requires_grad_(True), feat_sources=pre_src, fused_leaky_relu=0.2, **kw)


def test_mask_pack_and_host_expand(cuda_lib):
    """rmvd_mask_pack_u8 (device) + rmvd_host_expand_mask_async (host function on the stream) reproduce a 0 / 1 float32 tensor bit for
    bit, for sizes with a tail and for an unaligned host destination."""
    import ctypes
    from robustmvd_b200 import _lib
    g = torch.Generator().manual_seed(5)
    for n in (4, 7, 1023, 70001, 3 * 65536 + 9):
        m = (torch.rand(n, generator=g) < 0.4).float().cuda()
        packed = torch.full((n + 4,), 9, dtype=torch.uint8, device="cuda")
        _lib.check(cuda_lib.rmvd_mask_pack_u8(_lib.ptr(m), _lib.ptr(packed), n, _lib.stream_ptr()), "pack")
        stage = torch.empty(n + 4, dtype=torch.uint8).pin_memory()
        stage.copy_(packed, non_blocking=True)
        host = torch.full((n + 9,), 7.0).pin_memory()
        dst = host[1:1 + n]                       # 4-byte aligned only
        _lib.check(cuda_lib.rmvd_host_expand_mask_async(ctypes.c_void_p(stage.data_ptr()), ctypes.c_void_p(dst.data_ptr()), n, 4,
                                                        _lib.stream_ptr()), "expand")
        torch.cuda.synchronize()
        assert torch.equal(dst, m.cpu()) and float(host[0]) == 7.0 and bool((host[1 + n:] == 7.0).all())
        assert bool((packed[n:] == 9).all())
        # one bit per element
        words = (n + 31) // 32
        pb = torch.full((words + 1,), -1, dtype=torch.int32, device="cuda")
        _lib.check(cuda_lib.rmvd_mask_pack_bits(_lib.ptr(m), _lib.ptr(pb), n, _lib.stream_ptr()), "pack bits")
        sb = torch.empty(words + 1, dtype=torch.int32).pin_memory()
        sb.copy_(pb, non_blocking=True)
        host.fill_(7.0)
        _lib.check(cuda_lib.rmvd_host_expand_mask_bits_async(ctypes.c_void_p(sb.data_ptr()), ctypes.c_void_p(dst.data_ptr()), n, 4,
                                                             _lib.stream_ptr()), "expand bits")
        torch.cuda.synchronize()
        assert torch.equal(dst, m.cpu()) and float(host[0]) == 7.0 and bool((host[1 + n:] == 7.0).all()) and int(pb[words]) == -1


def test_cfg4_full_size_against_reference_golden(cuda_lib):
    """BASELINE.json configs[3] at full size (ETH3D 96x144 features, C=256, 256 planes, 2 source views) against the golden frozen
    from the unmodified reference (tools/make_fullsize_golden.py): masks, geometric masks and corner indices bit for bit, the
    correlation on a stride-31 subsample within 1e-4 of max|ref|."""
    case = cases.fullsize_planesweep_case()
    g = load_golden("ps", case["name"])
    corrs, masks, inv, dbg = _run_cuda(case)
    check_planesweep_against_golden(g, corrs, masks, inv, dbg["ix0"], dbg["iy0"], dbg["geom"], tol=TOL)
    corrs2, masks2, _, _ = _run_cuda(case, debug=False)           # the production kernel variant (no debug outputs)
    for v in range(len(corrs)):
        assert np.array_equal(masks2[v], masks[v]) and np.array_equal(corrs2[v], corrs[v])


@pytest.mark.parametrize("name", ["roll_pitch_12x20", "src_res_differs", "behind_12x20", "stereo_x_12x20", "explicit_inv_rank4"])
def test_block_state_like_the_reference(cuda_lib, name):
    """The reference block leaves ``coeffs`` and ``sampling_points`` on itself (planesweep_corr.py:389-394, :489-512); here they
    are materialised on first read (``rmvd_sampling_points_fwd``): us / vs / zs and the geometric mask bit for bit against the
    oracle's chain, the derived coefficient properties against their definitions."""
    from robustmvd_b200.blocks.planesweep_corr import PlanesweepCorrelation
    case = [c for c in PS if c["name"] == name][0]
    t = lambda a: None if a is None else torch.from_numpy(np.asarray(a)).cuda()  # noqa: E731
    kw = {k: (t(v) if isinstance(v, np.ndarray) else v) for k, v in case["kwargs"].items()}
    blk = PlanesweepCorrelation()
    blk(feat_key=t(case["feat_key"]), intrinsics_key=t(case["intrinsics_key"]), feat_sources=[t(f) for f in case["feat_sources"]],
        source_to_key_transforms=[t(x) for x in case["transforms"]],
        intrinsics_sources=None if case["intrinsics_sources"] is None else [t(x) for x in case["intrinsics_sources"]], **kw)
    _, _, _, odbg = oracle.planesweep_correlation(case["feat_key"], case["intrinsics_key"], case["feat_sources"], case["transforms"],
                                                  case["intrinsics_sources"], return_debug=True, **case["kwargs"])
    assert len(blk.sampling_points) == len(blk.coeffs) == len(case["feat_sources"])
    for v, sp in enumerate(blk.sampling_points):
        shape = sp.us.shape
        assert np.array_equal(sp.us.cpu().numpy(), np.broadcast_to(odbg[v]["us"], shape))
        assert np.array_equal(sp.vs.cpu().numpy(), np.broadcast_to(odbg[v]["vs"], shape))
        assert np.array_equal(sp.zs.cpu().numpy(), np.broadcast_to(odbg[v]["zs"], shape), equal_nan=True)
        assert sp.mask.dtype == torch.bool and np.array_equal(sp.mask.cpu().numpy(), np.broadcast_to(odbg[v]["geom_mask"], shape))
        assert sp.get_uvs().shape == (shape[0], shape[1], 2, shape[2], shape[3]) and sp.num == shape[1]
        co = blk.coeffs[v]
        N, _, h, w = case["feat_key"].shape
        assert co.u_infs_h.shape == (N, 1, h, w) and co.m_k_h.shape == (N, 1, 1, 1) and (co.height, co.width) == (h, w)
        # us = (u_infs_h + m_u_h * d) / (k_infs_h + m_k_h * d) (:333-335) reproduces the sampling points from the coefficients
        d = 1.0 / sp.zs
        us2 = (co.u_infs_h + co.m_u_h * d) / (co.k_infs_h + co.m_k_h * d)
        ok = torch.isfinite(us2) & (sp.us.abs() < 1e8)
        assert torch.allclose(us2[ok], sp.us[ok], rtol=1e-4, atol=1e-3)
        assert torch.isfinite(co.u_infs).all() and co.z_poles.shape == (N, 1, h, w) and len(co.epipole) == 2


def test_torchcorr_border_padding(cuda_lib):
    """TorchCorr(padding_mode="border") through rmvd_corr_grid_fwd against the golden of the unmodified reference and the oracle."""
    import os
    from robustmvd_b200.blocks.planesweep_corr import TorchCorr
    c = cases.border_case()
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "border_torchcorr.npz"))
    cu = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
    for nm in ("dim", False, True):
        for wm in (0, 1):
            corr, mask = TorchCorr(normalize=nm, padding_mode="border")(cu(c["feat_key"]), cu(c["feat_src"]), cu(c["grids"]),
                                                                       cu(c["geom"]) if wm else None)
            assert np.array_equal(mask.cpu().numpy().astype(np.uint8), g[f"mask_{nm}_{wm}"])
            ref = g[f"corr_{nm}_{wm}"]
            assert np.abs(corr.cpu().numpy() - ref).max() <= TOL * np.abs(ref).max()


@pytest.mark.parametrize("padding", ["zeros", "border"])
def test_warp_multi_and_warponlycorr_one_launch(cuda_lib, padding):
    """warp_multi (planesweep_corr.py:13-46) and WarpOnlyCorr.forward (:107-140) on caller grids / offsets: one launch for all S
    planes, against the reference's own formulation restated with torch CPU ops (per-plane warp(): grid scaling :85-88, grid_sample,
    mask thresholds :95-102)."""
    import torch.nn.functional as F
    from robustmvd_b200.blocks.planesweep_corr import warp_multi, WarpOnlyCorr
    rng = np.random.default_rng(11)
    N, C, hs, ws, S, h, w = 2, 6, 9, 14, 5, 7, 8
    x = torch.from_numpy(rng.standard_normal((N, C, hs, ws), dtype=np.float32))
    grids = torch.from_numpy(np.stack([rng.uniform(-3, ws + 3, (N, S, h, w)), rng.uniform(-3, hs + 3, (N, S, h, w))], 2).astype(np.float32))

    def ref_warp(g2):                                            # g2 [N,2,h,w] in source pixels
        gg = g2.permute(0, 2, 3, 1)
        gn = torch.cat([2 * gg[..., :1] / ws - 1, 2 * gg[..., 1:] / hs - 1], -1)
        out = F.grid_sample(x, gn, padding_mode=padding, align_corners=False)
        if padding == "border":
            return out, torch.ones((N, 1, h, w))
        m = F.grid_sample(torch.ones((N, 1, hs, ws)), gn, padding_mode="zeros", align_corners=False)
        m[m < 0.9999] = 0
        m[m > 0] = 1
        return out, m
    ro, rm = zip(*[ref_warp(grids[:, s]) for s in range(S)])
    ro, rm = torch.stack(ro, 1), torch.stack([m[:, 0] for m in rm], 1)
    out, mask = warp_multi(x.cuda(), grids=grids.cuda(), padding_mode=padding)
    assert out.shape == (N, S, C, h, w) and mask.shape == (N, S, h, w)
    assert torch.equal(mask.cpu(), rm) and float((out.cpu() - ro).abs().max()) < 1e-5
    # offsets relative to the pixel centres (:68-79)
    yy, xx = torch.meshgrid(torch.arange(h), torch.arange(w), indexing="ij")
    centres = torch.stack((xx + 0.5, yy + 0.5), 0).float().view(1, 1, 2, h, w)
    out2, mask2 = warp_multi(x.cuda(), offsets=(grids - centres).cuda(), padding_mode=padding)
    assert float((out2.cpu() - ro).abs().max()) < 1e-4 and float((mask2.cpu() != rm).float().mean()) < 0.01
    # WarpOnlyCorr: normalisation after the warp (:133-134), then masking (:136)
    wo, wm = WarpOnlyCorr(normalize="after", padding_mode=padding)(None, x.cuda(), grids.cuda())
    expect = ro / (torch.linalg.norm(ro, dim=2, keepdim=True) + 1e-9) * rm.unsqueeze(2)
    assert float((wo.cpu() - expect).abs().max()) < 1e-5 and torch.equal(wm.cpu(), rm)


def test_graphed_planesweep_replays_the_block(cuda_lib):
    """pipeline.GraphedPlanesweep: the block's launch sequence captured in a CUDA graph; replays with new inputs of the captured
    shape equal the eager call bit for bit."""
    from robustmvd_b200.blocks.planesweep_corr import PlanesweepCorrelation
    from robustmvd_b200.pipeline import GraphedPlanesweep
    cu = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731

    def kwargs(seed):
        inp = syn.planesweep_inputs(24, 40, 64, 2, seed=seed)
        return dict(feat_key=cu(inp["feat_key"]), intrinsics_key=cu(inp["intrinsics_key"]), feat_sources=[cu(f) for f in inp["feat_sources"]],
                    source_to_key_transforms=[cu(t) for t in inp["transforms"]], num_sampling_points=48, min_depth=0.4, max_depth=1000.0)
    gp = GraphedPlanesweep(kwargs(1))
    blk = PlanesweepCorrelation()
    for seed in (1, 2, 3):
        kw = kwargs(seed)
        c0, m0, i0 = blk(**kw)
        c1, m1, i1 = gp(**kw)
        torch.cuda.synchronize()
        assert all(torch.equal(a, b) for a, b in zip(c0, c1)) and all(torch.equal(a, b) for a, b in zip(m0, m1)) and torch.equal(i0, i1)
    with pytest.raises(ValueError):
        gp(**dict(kwargs(1), num_sampling_points=32))
