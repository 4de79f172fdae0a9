"""Per-kernel timing of every entry point of the path at the BASELINE.json shapes (1 GPU), against the HBM roofline.

    python tools/bench_kernels.py [--reps 20] [--only NAME]

Each kernel is timed with CUDA events on the launching stream, one event pair per repetition, with a 256 MB scratch
write between repetitions so that nothing is served from the 126 MB L2 (cold = `ms_cold`; `ms_warm` = back-to-back
without the flush; `ms_warm_graph` = the same loop replayed from a CUDA graph, i.e. without host time between launches).  `frac` = algorithmic bytes / ms_cold / MEASURED_PEAKS.json hbm_gbs.  One JSON line per kernel.
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from robustmvd_b200 import synthetic as syn  # noqa: E402
from robustmvd_b200.blocks import utils as bu  # noqa: E402
from robustmvd_b200.blocks.planesweep_corr import PlanesweepCorrelation  # noqa: E402

cu = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731


def time_fn(fn, reps, flush):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    cold = []
    for _ in range(reps):
        flush.fill_(1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        cold.append(a.elapsed_time(b))
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    warm = a.elapsed_time(b) / reps
    # the same back-to-back loop replayed from a CUDA graph: no Python, ctypes or allocator time between the launches (for kernels of
    # a few tens of microseconds `ms_warm` is the host's call rate, not the kernel)
    graph_ms = None
    try:
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            fn()
        torch.cuda.current_stream().wait_stream(side)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(reps):
                fn()
        g.replay()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        g.replay()
        b.record()
        torch.cuda.synchronize()
        graph_ms = a.elapsed_time(b) / reps
    except Exception:  # noqa: BLE001  (a wrapper that synchronises cannot be captured: the number is simply left out)
        torch.cuda.synchronize()
    return float(np.mean(cold)), float(np.min(cold)), warm, graph_ms


def psweep_case(cfg, V, frames, S=None):
    S = S or cfg["S"]
    inp = syn.planesweep_inputs(cfg["h"], cfg["w"], cfg["C"], V, n=frames)
    kw = dict(feat_key=cu(inp["feat_key"]), intrinsics_key=cu(inp["intrinsics_key"]),
              feat_sources=[cu(f) for f in inp["feat_sources"]], source_to_key_transforms=[cu(t) for t in inp["transforms"]],
              num_sampling_points=S, min_depth=cfg["min_depth"], max_depth=cfg["max_depth"])
    P = cfg["h"] * cfg["w"]
    nbytes = 4 * frames * cfg["C"] * P * (1 + V) + 8 * frames * V * S * P
    return kw, nbytes, frames * V * S * P


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--only", default=None)
    args = ap.parse_args()
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) \
        else {"hbm_gbs": 6650.0}
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    jobs = []

    def add(name, fn, nbytes, units, unit_name):
        if args.only is None or args.only in name:
            jobs.append((name, fn, nbytes, units, unit_name))

    blk = PlanesweepCorrelation()
    for label, cfg, V, frames in (("cfg2_kitti_1frame", syn.CFG2, 2, 1), ("cfg2_kitti_8frames", syn.CFG2, 2, 8),
                                  ("cfg4_eth3d_1frame_v2", syn.CFG4, 2, 1), ("cfg4_eth3d_1frame_v10", syn.CFG4, 10, 1)):
        kw, nb, vox = psweep_case(cfg, V, frames)
        add(f"psweep_corr_{label}", (lambda kw=kw: blk(**kw)), nb, vox, "voxels")
    # backward of the correlation at cfg2, one frame: reads key + sources + grad_corr, writes their gradients
    kwb, _, voxb = psweep_case(syn.CFG2, 2, 1)
    kwb["feat_key"].requires_grad_(True)
    for f in kwb["feat_sources"]:
        f.requires_grad_(True)
    bcorrs = blk(**kwb)[0]
    bg = [torch.randn_like(x) for x in bcorrs]
    bin_ = [kwb["feat_key"]] + kwb["feat_sources"]
    add("psweep_corr_bwd_cfg2_kitti_1frame", lambda: torch.autograd.grad(bcorrs, bin_, bg, retain_graph=True),
        2 * 4 * 256 * 7680 * 3 + 4 * voxb, voxb, "voxels")
    kwb8, _, voxb8 = psweep_case(syn.CFG2, 2, 8)
    kwb8["feat_key"].requires_grad_(True)
    for f in kwb8["feat_sources"]:
        f.requires_grad_(True)
    bcorrs8 = blk(**kwb8)[0]
    bg8 = [torch.randn_like(x) for x in bcorrs8]
    bin8 = [kwb8["feat_key"]] + kwb8["feat_sources"]
    add("psweep_corr_bwd_cfg2_kitti_8frames", lambda: torch.autograd.grad(bcorrs8, bin8, bg8, retain_graph=True),
        8 * (2 * 4 * 256 * 7680 * 3) + 4 * voxb8, voxb8, "voxels")
    c3 = syn.CFG3
    m = syn.mvs_inputs(c3["h"], c3["w"], c3["C"], c3["V"], c3["D"])
    feats, proj, depth, cost = cu(m["feats"]), cu(m["proj"]), cu(m["depth"]), cu(m["cost"])
    P3 = c3["h"] * c3["w"]
    add("mvs_variance_cfg3", lambda: bu.mvs_variance_volume(feats, proj, depth),
        4 * c3["C"] * P3 * c3["V"] + 4 * c3["C"] * c3["D"] * P3, c3["D"] * P3, "voxels")
    add("mvs_regress_cfg3", lambda: bu.mvs_regress(cost, depth), 4 * c3["D"] * P3 + 8 * P3, c3["D"] * P3, "voxels")
    # training direction of the MVSNet path (rmvd_mvs_variance_bwd / rmvd_mvs_regress_bwd) at cfg3
    featsg = feats.clone().requires_grad_(True)
    varg = bu.mvs_variance_volume(featsg, proj, depth)
    gvar = torch.randn_like(varg)
    add("mvs_variance_bwd_cfg3", lambda: torch.autograd.grad(varg, featsg, gvar, retain_graph=True),
        2 * 4 * c3["C"] * P3 * c3["V"] + 4 * c3["C"] * c3["D"] * P3, c3["D"] * P3, "voxels")
    costg = cost.clone().requires_grad_(True)
    depg = bu.mvs_regress(costg, depth)["depth"]
    gdep = torch.randn_like(depg)
    add("mvs_regress_bwd_cfg3", lambda: torch.autograd.grad(depg, costg, gdep, retain_graph=True), 8 * c3["D"] * P3 + 8 * P3, c3["D"] * P3, "voxels")
    add("homo_warp_cfg3", lambda: bu.homo_warp(feats[:, 1], proj[:, 1], proj[:, 0], depth),
        4 * c3["C"] * P3 + 4 * c3["C"] * c3["D"] * P3, c3["D"] * P3, "voxels")
    add("soft_argmin_cfg3", lambda: bu.soft_argmin(cost, 1), 8 * c3["D"] * P3, c3["D"] * P3, "voxels")
    wblk = PlanesweepCorrelation(warp_only=True, normalize=False)
    kww, _, _ = psweep_case(syn.CFG2, 1, 1, S=64)
    add("psweep_warp_only_cfg2_S64", lambda: wblk(**kww), 4 * 256 * 7680 + 4 * 64 * 257 * 7680, 64 * 7680, "voxels")
    # learned view fusion, elementwise part at cfg2 x 8 frames: reads corr + mask of V views, writes fused corr + mask
    from robustmvd_b200 import _lib
    Vf, Nf, Sf, Pf = 2, 8, 256, 48 * 160
    fl = [torch.randn((Nf, 1, 48, 160), device="cuda") for _ in range(Vf)]
    fc = [torch.randn((Nf, Sf, 48, 160), device="cuda") for _ in range(Vf)]
    fm = [(torch.rand((Nf, Sf, 48, 160), device="cuda") < 0.6).float() for _ in range(Vf)]
    fw = torch.empty((Vf, Nf, 1, 48, 160), device="cuda")
    fo = [torch.empty((Nf, Sf, 48, 160), device="cuda") for _ in range(2)]

    def run_fusion():
        _lib.check(_lib.load().rmvd_view_fusion_fwd(_lib.ptr_array(fl), _lib.ptr_array(fc), _lib.ptr_array(fm), Vf, Nf, Sf, 48, 160,
                                                    _lib.ptr(fw), _lib.ptr(fo[0]), _lib.ptr(fo[1]), _lib.stream_ptr()), "fusion")
    add("view_fusion_cfg2_8frames", run_fusion, 4 * (2 * Vf + 2) * Nf * Sf * Pf + 8 * Vf * Nf * Pf, Nf * Sf * Pf, "voxels")
    # Vis-MVSNet fused homography warp + group-wise correlation: 128x160 features, C=32, 8 groups, 64 planes
    vn, vc, vh, vw, vd, vg = 1, 32, 128, 160, 64, 8
    vref, vsrc = torch.randn((vn, vc, vh, vw), device="cuda"), torch.randn((vn, vc, vh, vw), device="cuda")
    vH = torch.eye(3, device="cuda").view(1, 1, 1, 1, 3, 3).repeat(vn, vd, 1, 1, 1, 1).contiguous()
    vH[:, :, 0, 0, 0, 2] = torch.linspace(2.0, 40.0, vd, device="cuda")       # plane-dependent disparity
    add("vis_groupcorr_128x160_c32_d64", lambda: bu.groupwise_cost_volume(vref, vsrc, vH, vg),
        4 * vn * vc * vh * vw * 2 + 4 * vn * vg * vd * vh * vw, vn * vd * vh * vw, "voxels")
    vcamL = torch.zeros((vn, 2, 4, 4), device="cuda")
    vcamL[:, 0] = torch.eye(4, device="cuda")
    vcamL[:, 1, :3, :3] = torch.tensor([[140.0, 0, 80.0], [0, 140.0, 64.0], [0, 0, 1.0]], device="cuda")
    vcamR = vcamL.clone()
    vcamR[:, 0, 0, 3] = 0.12
    vds = (0.45 + 0.05 * torch.rand((vn, 1, vh, vw), device="cuda")).contiguous()      # per-pixel schedule (refinement stage)
    vdi = torch.full((vn, 1, vh, vw), 0.008, device="cuda")
    add("vis_cost_volume_from_cameras_perpixel_128x160_c32_d64",
        lambda: bu.vis_cost_volume(vref, vsrc, vcamL, vcamR, vd, vds, vdi, vg),
        4 * vn * vc * vh * vw * 2 + 8 * vn * vh * vw + 4 * vn * vg * vd * vh * vw, vn * vd * vh * vw, "voxels")
    # CVP-MVSNet proj_cost: refinement level 128x160, C=16, 2 source views, 8 per-pixel hypotheses
    import types
    from robustmvd_b200.blocks import cvp_mvsnet_components as cvpc
    cB, cC, cH, cW, cD, cV = 1, 16, 128, 160, 8, 2
    cref = torch.randn((cB, cC, cH, cW), device="cuda")
    csrc = [[torch.randn((cB, cC, cH, cW), device="cuda")] for _ in range(cV)]
    cK = torch.tensor([[140.0, 0, 80.0], [0, 140.0, 64.0], [0, 0, 1.0]], device="cuda").repeat(cB, 1, 1)
    cE = torch.eye(4, device="cuda").repeat(cB, 1, 1)
    cEs = cE[:, None].repeat(1, cV, 1, 1).contiguous()
    cEs[:, :, 0, 3] = torch.tensor([0.1, -0.1], device="cuda")
    chyp = (0.6 + 0.02 * torch.arange(cD, device="cuda").view(1, cD, 1, 1) + 0.1 * torch.rand((cB, 1, cH, cW), device="cuda")).contiguous()
    cset = types.SimpleNamespace(nsrc=cV, mode="test")
    add("cvp_proj_cost_128x160_c16_d8", lambda: cvpc.proj_cost(cset, cref, csrc, 0, cK, cK[:, None].repeat(1, cV, 1, 1), cE, cEs, chyp),
        4 * cC * cH * cW * (1 + cV) + 4 * cD * cH * cW + 4 * cC * cD * cH * cW, cD * cH * cW, "voxels")
    for name, fn, nbytes, units, unit_name in jobs:
        cold, best, warm, graph_ms = time_fn(fn, args.reps, flush)
        print(json.dumps({"kernel": name, "ms_cold": cold, "ms_cold_min": best, "ms_warm": warm, "ms_warm_graph": graph_ms, "algorithmic_bytes": nbytes,
                          "achieved_gbs": nbytes / cold / 1e6, "frac_of_measured_hbm": nbytes / cold / 1e6 / peaks["hbm_gbs"],
                          f"{unit_name}_per_s": units / (cold * 1e-3)}), flush=True)


if __name__ == "__main__":
    main()
