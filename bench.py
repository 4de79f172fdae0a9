#!/usr/bin/env python
"""Benchmark of the plane-sweep cost-volume hot path (BASELINE.json metric: cost-volume voxels/s @384x1280).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--frames F] [--no-secondary]

Headline workload (configs[1]): robust_mvd plane-sweep correlation at the KITTI shape 384x1280 -> 48x160 features, C = 256,
S = 256 planes, 2 source views.  One step = one call of PlanesweepCorrelation.forward on a batch of F frames per GPU
(default 16: 377 MB of features in, 503 MB of corr + mask out, i.e. larger than the 126 MB L2, so consecutive steps
cannot be served from cache).  N > 1: one process per GPU (torchrun), each rank sweeps its own frames (weak scaling
by samples, no data-path collective); the time is the max over ranks.

Prints ONE JSON line (see the round contract): value = voxels/s with inputs resident in HBM; e2e = the same through
the block API from pinned HOST buffers with the H2D / D2H copies inside the timed region; roofline = algorithmic
bytes / device time against MEASURED_PEAKS.json; cpu_baseline = the reference's own files on the host cores
(oracle/_ref, kind "reference") or, where they are absent, oracle/torch_cpu.py (kind "port"), on a bounded sample.
The other BASELINE.json configs ride on the same line:

  secondary          configs[1] at one frame (the reference's eval batch size), configs[3] shape at one frame (V = 2, 10),
                     configs[2] MVSNet variance volume + regression (each with roofline, cpu_baseline, e2e)
  torch_gpu_baseline the reference's operator sequence (oracle/torch_cpu.py) on stock torch CUDA ops on the same B200, and the
                     mask / index agreement of our kernel with CUDA-ATen grid_sample
  e2e_model          model.run(**sample) of robust_mvd from host images to host depth (configs[0] shape of work)
  sweep              configs[4]: planes x views grid at the configs[1] spatial shape, samples sharded over the ranks
  slab               (N > 1) configs[3]: depth slabs over the ranks with the fused peer-store gather
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

from robustmvd_b200 import synthetic as syn  # noqa: E402

CFG = syn.CFG2
KW = dict(num_sampling_points=CFG["S"], min_depth=CFG["min_depth"], max_depth=CFG["max_depth"])
METRIC = "planesweep_cost_volume_voxels_per_s"


def voxels(frames, cfg=CFG, V=None, S=None):
    return frames * (V or cfg["V"]) * (S or cfg["S"]) * cfg["h"] * cfg["w"]


def algorithmic_bytes(frames, cfg=CFG, V=None, S=None):
    """features read once, corr + mask written once."""
    P = cfg["h"] * cfg["w"]
    V, S = V or cfg["V"], S or cfg["S"]
    return 4 * frames * cfg["C"] * P * (1 + V) + 2 * 4 * frames * V * S * P


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)), "measured"
    return {"hbm_gbs": 6650.0}, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.index), "-lms", "20"], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for k, n in enumerate(names) if any(len(r) > 2 + k and r[2 + k] == "Active" for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def make_inputs(frames, seed, cfg=CFG, V=None):
    return syn.planesweep_inputs(cfg["h"], cfg["w"], cfg["C"], V or cfg["V"], n=frames, seed=seed)


# ------------------------------------------------------------------------------------------------------------------
# CPU baselines (the only places where bench.py executes anything under oracle/)
# ------------------------------------------------------------------------------------------------------------------
def _reference_block():
    """The reference's own PlanesweepCorrelation (unmodified file under /root/reference or oracle/_ref), or None."""
    try:
        from tools import ref_loader
        if ref_loader.available():
            return ref_loader.planesweep_corr().PlanesweepCorrelation()
    except Exception:  # noqa: BLE001
        pass
    return None


def cpu_reference_step(inp1, block=None):
    import torch
    if block is not None:
        t = lambda a: torch.from_numpy(np.ascontiguousarray(a))  # noqa: E731
        with torch.no_grad():
            return block(feat_key=t(inp1["feat_key"]), intrinsics_key=t(inp1["intrinsics_key"]),
                         feat_sources=[t(f) for f in inp1["feat_sources"]],
                         source_to_key_transforms=[t(x) for x in inp1["transforms"]], **KW)
    from oracle import torch_cpu
    return torch_cpu.planesweep_correlation(inp1["feat_key"], inp1["intrinsics_key"], inp1["feat_sources"],
                                            inp1["transforms"], **KW)


def time_cpu_baseline(steps, warmup):
    """The reference's algorithm on the host cores, one frame per step (bounded sample of the workload)."""
    import torch
    torch.set_num_threads(os.cpu_count())
    inp1 = make_inputs(1, 1234)
    block = _reference_block()
    for _ in range(warmup):
        cpu_reference_step(inp1, block)
    t0 = time.perf_counter()
    for _ in range(steps):
        cpu_reference_step(inp1, block)
    dt = (time.perf_counter() - t0) / steps
    what = ("the reference's own planesweep_corr.py (unmodified, oracle/_ref or /root/reference) on torch CPU" if block is not None
            else "oracle/torch_cpu.py = the reference's matmul + grid_sample sequence on torch CPU")
    return dict(value=voxels(1) / dt, unit="voxels/s", cores=torch.get_num_threads(), kind="reference" if block is not None else "port",
                sample=f"1 frame of the workload per step ({steps} steps, {dt * 1e3:.0f} ms/frame); {what}", ms_per_frame=dt * 1e3)


def time_cpu_mvs(planes=24):
    """MVSNet variance volume (a bounded slab of `planes` of the 192 planes) and regression (full size) on the host cores."""
    import torch
    from oracle import torch_cpu
    torch.set_num_threads(os.cpu_count())
    c3 = syn.CFG3
    m = syn.mvs_inputs(c3["h"], c3["w"], c3["C"], c3["V"], c3["D"])
    t0 = time.perf_counter()
    torch_cpu.mvs_variance(m["feats"], m["proj"], m["depth"][:, :planes])
    dt_var = time.perf_counter() - t0
    t0 = time.perf_counter()
    torch_cpu.mvs_regress(m["cost"], m["depth"])
    dt_reg = time.perf_counter() - t0
    P = c3["h"] * c3["w"]
    return (dict(value=planes * P / dt_var, unit="voxels/s", cores=torch.get_num_threads(), kind="port",
                 sample=f"{planes} of {c3['D']} planes, all {c3['V']} views ({dt_var * 1e3:.0f} ms); oracle/torch_cpu.mvs_variance = "
                        "homo_warp + running sums of mvsnet.py:123-136 on torch CPU"),
            dict(value=P / dt_reg, unit="pixels/s", cores=torch.get_num_threads(), kind="port",
                 sample=f"full volume, {dt_reg * 1e3:.0f} ms; softmax + depth_regression + 4-bin confidence (mvsnet.py:140-160) on torch CPU"))


def base_line(args, frames, world):
    return {"metric": METRIC, "unit": "voxels/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "configs[1]: robust_mvd plane-sweep correlation, KITTI 384x1280 -> 48x160 features, "
                                   "C=256, S=256 planes (0.4 m..1000 m, linear inverse depth), 2 source views",
                       "frames_per_gpu_per_step": frames, "frames_per_step": frames * world,
                       "voxels_per_step": voxels(frames) * world, "parallelism": f"samples x{world}",
                       "l2": "step working set %d MB > 126 MB L2 (no flush needed)" % round(algorithmic_bytes(frames) / 1e6)
                             if frames >= 4 else
                             "L2 flushed between steps (256 MB scratch write)"}}


def run_reference(args):
    """--impl reference: the reference's CPU path for the same metric, rank 0 only."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    cb = time_cpu_baseline(max(1, args.steps), min(args.warmup, 1))
    line = base_line(args, 1, 1)
    line.update({"impl": "reference", "value": cb["value"], "ms_per_step": cb["ms_per_frame"], "n_gpus": args.gpus,
                 "frames_per_s": 1e3 / cb["ms_per_frame"], "cpu_baseline": cb, "gpu_launches": 0,
                 "e2e": {"value": cb["value"], "unit": "voxels/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}})
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------------------------
# secondary records (rank 0, after the headline measurement)
# ------------------------------------------------------------------------------------------------------------------
def _cold_warm(torch, fn, flush, reps=10):
    """ms per call: cold (256 MB scratch write before every call) and warm (back to back), CUDA events on the current stream."""
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
    return float(np.median(cold)), a.elapsed_time(b) / reps


def secondary_records(torch, dev, peaks, with_cpu):
    from robustmvd_b200.blocks import utils as bu
    from robustmvd_b200.blocks.planesweep_corr import PlanesweepCorrelation
    cu = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)  # noqa: E731
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    hbm = peaks["hbm_gbs"]
    out = []

    def psweep(name, cfg, V, frames=1):
        inp = make_inputs(frames, 4321, cfg, V)
        kw = dict(feat_key=cu(inp["feat_key"]), intrinsics_key=cu(inp["intrinsics_key"]), feat_sources=[cu(f) for f in inp["feat_sources"]],
                  source_to_key_transforms=[cu(t) for t in inp["transforms"]], num_sampling_points=cfg["S"], min_depth=cfg["min_depth"],
                  max_depth=cfg["max_depth"])
        blk = PlanesweepCorrelation()
        cold, warm = _cold_warm(torch, lambda: blk(**kw), flush)
        ab = algorithmic_bytes(frames, cfg, V)
        # the same call as a CUDA graph (robustmvd_b200.pipeline.GraphedPlanesweep): device time of the launch sequence without the host
        ms_graph = None
        try:
            from robustmvd_b200.pipeline import GraphedPlanesweep
            gp = GraphedPlanesweep(kw)
            c0, m0, _ = blk(**kw)
            c1, m1, _ = gp(**kw)
            same = all(torch.equal(a, b) for a, b in zip(c0, c1)) and all(torch.equal(a, b) for a, b in zip(m0, m1))
            ms_graph = dict(zip(("cold", "warm"), _cold_warm(torch, lambda: gp.graph.replay(), flush)))
            if not same:
                ms_graph = "graph replay differs from the eager call"
        except Exception as e:  # noqa: BLE001
            ms_graph = repr(e)[:200]
        out.append({"name": name, "metric": METRIC, "unit": "voxels/s", "value": voxels(frames, cfg, V) / (cold * 1e-3), "ms_cold": cold,
                    "ms_warm": warm, "ms_cuda_graph": ms_graph, "frames": frames, "views": V,
                    "roofline": {"bound": "hbm", "achieved": ab / cold / 1e6, "peak": hbm, "unit": "GB/s", "frac": ab / cold / 1e6 / hbm,
                                 "algorithmic_bytes_per_launch": ab, "timing": "ms_cold"}})

    psweep("configs[1] at one frame (the reference's evaluation batch size)", CFG, 2)
    psweep("configs[3] shape (ETH3D 768x1152 -> 96x144, C=256, S=256), one frame, 2 source views, one GPU", syn.CFG4, 2)
    psweep("configs[3] shape, one frame, 10 source views, one GPU", syn.CFG4, 10)

    # configs[2]: MVSNet variance volume + regression at the DTU shape
    c3 = syn.CFG3
    m = syn.mvs_inputs(c3["h"], c3["w"], c3["C"], c3["V"], c3["D"])
    feats, proj, depth, cost = cu(m["feats"]), cu(m["proj"]), cu(m["depth"]), cu(m["cost"])
    P = c3["h"] * c3["w"]
    cpu_var, cpu_reg = time_cpu_mvs() if with_cpu else (None, None)
    cold, warm = _cold_warm(torch, lambda: bu.mvs_variance_volume(feats, proj, depth), flush, reps=5)
    ab = 4 * c3["C"] * P * c3["V"] + 4 * c3["C"] * c3["D"] * P
    # stock PyTorch on the same B200 running the reference's operator sequence (the bar for every kernel)
    from oracle import torch_cpu

    def stock(fn, reps):
        try:
            with torch.no_grad():
                ref = fn()
                c, _ = _cold_warm(torch, fn, flush, reps=reps)
            return ref, c
        except Exception as e:  # noqa: BLE001
            return None, repr(e)[:200]
    ref_var, stock_var_ms = stock(lambda: torch_cpu.mvs_variance(feats, proj, depth, device=dev, return_torch=True), 3)
    stock_var = {"what": "oracle/torch_cpu.mvs_variance with device='cuda': grid_sample per view + the variance expression (mvsnet.py:123-136)",
                 "ms_cold": stock_var_ms}
    if ref_var is not None:
        ours = bu.mvs_variance_volume(feats, proj, depth)
        stock_var["max_abs_diff_over_max"] = float((ours - ref_var).abs().max() / ref_var.abs().max())
        del ours, ref_var
    # e2e: pinned host features in, variance volume back in pinned host memory
    h_feats = torch.from_numpy(m["feats"]).pin_memory()
    h_var = torch.empty((1, c3["C"], c3["D"], c3["h"], c3["w"]), dtype=torch.float32).pin_memory()

    def var_e2e():
        f = h_feats.to(dev, non_blocking=True)
        h_var.copy_(bu.mvs_variance_volume(f, proj, depth), non_blocking=True)
        torch.cuda.synchronize()
    var_e2e()
    t0 = time.perf_counter()
    for _ in range(3):
        var_e2e()
    ms_e2e = (time.perf_counter() - t0) / 3 * 1e3
    out.append({"name": "configs[2]: MVSNet variance cost volume, DTU 896x1216 -> 224x304, C=32, 5 views, 192 planes",
                "metric": "mvs_variance_voxels_per_s", "unit": "voxels/s", "value": c3["D"] * P / (cold * 1e-3), "ms_cold": cold, "ms_warm": warm,
                "roofline": {"bound": "hbm", "achieved": ab / cold / 1e6, "peak": hbm, "unit": "GB/s", "frac": ab / cold / 1e6 / hbm,
                             "algorithmic_bytes_per_launch": ab, "timing": "ms_cold"},
                "e2e": {"value": c3["D"] * P / (ms_e2e * 1e-3), "unit": "voxels/s", "ms_per_step": ms_e2e, "h2d_bytes_per_step": h_feats.numel() * 4,
                        "d2h_bytes_per_step": h_var.numel() * 4, "api": "blocks.utils.mvs_variance_volume from pinned host features, volume copied back"},
                "torch_gpu_baseline": stock_var, "cpu_baseline": cpu_var})
    cold, warm = _cold_warm(torch, lambda: bu.mvs_regress(cost, depth), flush)
    ab = 4 * c3["D"] * P + 8 * P
    ref_reg, stock_reg_ms = stock(lambda: torch_cpu.mvs_regress(cost, depth, device=dev, return_torch=True), 5)
    stock_reg = {"what": "oracle/torch_cpu.mvs_regress with device='cuda': softmax, two weighted sums, pad + avg_pool3d, gather (mvsnet.py:140-160)",
                 "ms_cold": stock_reg_ms}
    if ref_reg is not None:
        stock_reg["depth_max_rel_diff"] = float(((bu.mvs_regress(cost, depth)["depth"] - ref_reg[0]).abs() / ref_reg[0].abs()).max())
    h_cost = torch.from_numpy(m["cost"]).pin_memory()
    h_depth = torch.empty((1, c3["h"], c3["w"]), dtype=torch.float32).pin_memory()

    def reg_e2e():
        r = bu.mvs_regress(h_cost.to(dev, non_blocking=True), depth)
        h_depth.copy_(r["depth"], non_blocking=True)
        torch.cuda.synchronize()
    reg_e2e()
    t0 = time.perf_counter()
    for _ in range(5):
        reg_e2e()
    ms_e2e = (time.perf_counter() - t0) / 5 * 1e3
    out.append({"name": "configs[2]: softmax + expected depth + 4-bin confidence over the regularised cost [1,192,224,304]",
                "metric": "mvs_regress_pixels_per_s", "unit": "pixels/s", "value": P / (cold * 1e-3), "ms_cold": cold, "ms_warm": warm,
                "roofline": {"bound": "hbm", "achieved": ab / cold / 1e6, "peak": hbm, "unit": "GB/s", "frac": ab / cold / 1e6 / hbm,
                             "algorithmic_bytes_per_launch": ab, "timing": "ms_cold"},
                "e2e": {"value": P / (ms_e2e * 1e-3), "unit": "pixels/s", "ms_per_step": ms_e2e, "h2d_bytes_per_step": h_cost.numel() * 4,
                        "d2h_bytes_per_step": h_depth.numel() * 4, "api": "blocks.utils.mvs_regress from a pinned host cost volume, depth copied back"},
                "torch_gpu_baseline": stock_reg, "cpu_baseline": cpu_reg})
    return out


def torch_gpu_baseline(torch, dev):
    """The reference's operator sequence on stock torch CUDA ops (same B200), and our kernel's agreement with it."""
    from oracle import torch_cpu
    from robustmvd_b200.blocks.planesweep_corr import PlanesweepCorrelation
    inp = make_inputs(1, 1234)
    cu = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)  # noqa: E731
    fk, fs = cu(inp["feat_key"]), [cu(f) for f in inp["feat_sources"]]

    def ref():
        return torch_cpu.planesweep_correlation(fk, inp["intrinsics_key"], fs, inp["transforms"], device=dev, return_torch=True, **KW)
    for _ in range(2):
        rc, rm, _ = ref()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    reps = 5
    for _ in range(reps):
        ref()
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) / reps * 1e3
    corrs, masks, _ = PlanesweepCorrelation()(feat_key=fk, intrinsics_key=cu(inp["intrinsics_key"]), feat_sources=fs,
                                              source_to_key_transforms=[cu(t) for t in inp["transforms"]], **KW)
    mism = sum(int((m != r).sum().item()) for m, r in zip(masks, rm))
    # corr compared where both masks agree (a flipped mask zeroes one side)
    err = max(float(((c - r).abs() * (m == q)).max().item() / max(float(r.abs().max().item()), 1e-30)) for c, r, m, q in zip(corrs, rc, masks, rm))
    return {"what": "oracle/torch_cpu.py's operator sequence (all-pairs matmul + 2x grid_sample + thresholds, planesweep_corr.py:152-195) "
                    "with device='cuda': stock PyTorch on the same B200, configs[1], one frame per step",
            "value": voxels(1) / (ms * 1e-3), "unit": "voxels/s", "ms_per_frame": ms, "timing": "wall clock around 5 synchronised calls",
            "mask_mismatches_vs_cuda_aten": mism, "mask_elements": int(sum(m.numel() for m in masks)), "corr_max_rel_err_vs_cuda_aten": err}


def e2e_model_record(torch, dev, reps=5):
    """model.run(**sample): host images -> host depth through the reference's model API (robust_mvd, random init)."""
    import robustmvd_b200 as rmvd
    torch.manual_seed(0)
    model = rmvd.create_model("robust_mvd", pretrained=False, weights=None, num_gpus=1)
    sample = syn.model_sample(384, 1280, 2)
    pred, _ = model.run(**sample)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        pred, _ = model.run(**sample)
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) / reps * 1e3
    nin = sum(int(np.asarray(im).nbytes) for im in sample["images"])
    return {"what": "robustmvd_b200.create_model('robust_mvd', pretrained=False).run(**sample): 3 host images 384x1280 (key + 2 sources) "
                    "-> host depth map; random-init weights, encoder / decoder in cuDNN, correlation + fusion in this library",
            "value": 1e3 / ms, "unit": "frames/s", "ms_per_frame": ms, "h2d_bytes_per_step": nin, "d2h_bytes_per_step": int(pred["depth"].nbytes),
            "timing": "wall clock around %d synchronised calls" % reps}


def sweep_records(torch, dist, dev, world, peaks, frames=4):
    """configs[4]: S in {64,128,256,512} x V in {2,4,6,8,10} at the configs[1] spatial shape, `frames` samples per rank and step."""
    from robustmvd_b200.blocks.planesweep_corr import PlanesweepCorrelation
    cu = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)  # noqa: E731
    blk = PlanesweepCorrelation()
    cells = []
    for V in (2, 4, 6, 8, 10):
        inp = make_inputs(frames, 99 + int(os.environ.get("RANK", "0")), CFG, V)
        fk, Kk = cu(inp["feat_key"]), cu(inp["intrinsics_key"])
        fs, Ts = [cu(f) for f in inp["feat_sources"]], [cu(t) for t in inp["transforms"]]
        for S in (64, 128, 256, 512):
            kw = dict(feat_key=fk, intrinsics_key=Kk, feat_sources=fs, source_to_key_transforms=Ts, num_sampling_points=S,
                      min_depth=CFG["min_depth"], max_depth=CFG["max_depth"])
            for _ in range(2):
                blk(**kw)
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            reps = 6
            e0.record()
            for _ in range(reps):
                blk(**kw)
            e1.record()
            torch.cuda.synchronize()
            ms = torch.tensor([e0.elapsed_time(e1) / reps], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(ms, op=dist.ReduceOp.MAX)
            ms = float(ms.item())
            ab = algorithmic_bytes(frames, CFG, V, S)
            cells.append({"S": S, "V": V, "ms_per_step": ms, "voxels_per_s": voxels(frames, CFG, V, S) * world / (ms * 1e-3),
                          "frac_of_hbm": ab / ms / 1e6 / peaks["hbm_gbs"]})
    return {"what": "configs[4]: planes x source views at 384x1280 -> 48x160, C=256; %d samples per GPU and step, samples sharded over %d GPU(s), "
                    "no collective; time = max over ranks; back-to-back steps (working set of the small cells fits L2)" % (frames, world),
            "cells": cells}


def slab_records(torch, dist, dev, world, rank):
    """configs[3]: depth slabs over the ranks, corr / mask gathered by peer stores from inside the plane-sweep kernel."""
    from robustmvd_b200 import parallel
    from robustmvd_b200.blocks.planesweep_corr import PlanesweepCorrelation
    from robustmvd_b200.blocks import utils as bu
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

    recs = []
    for V in (2, 10):
        inp = syn.planesweep_inputs(cfg["h"], cfg["w"], cfg["C"], V, seed=77)      # same seed on every rank: replicated features
        kw = dict(feat_key=cu(inp["feat_key"]), intrinsics_key=cu(inp["intrinsics_key"]), feat_sources=[cu(f) for f in inp["feat_sources"]],
                  source_to_key_transforms=[cu(t) for t in inp["transforms"]], num_sampling_points=cfg["S"], min_depth=cfg["min_depth"],
                  max_depth=cfg["max_depth"])
        blk = PlanesweepCorrelation()
        full_c, full_m, _ = blk(**kw)
        ms_full = timed(lambda: blk(**kw))
        ms_slab = timed(lambda: blk(plane_range=parallel.slab_range(cfg["S"], world, rank), **kw))
        out_bytes = 8 * V * cfg["S"] * cfg["h"] * cfg["w"]
        rec = {"views": V, "frames": 1, "ms_1gpu": ms_full, "ms_plane_slab_compute_only": ms_slab,
               "nvlink_bytes_received_per_rank": out_bytes * (world - 1) // world,
               "nvlink_floor_ms_at_770GBs": out_bytes * (world - 1) / world / 770e9 * 1e3}
        # "planes" / "tiles": the fused kernel, the ranks split the planes (depth slabs proper) or the tiles; every rank's sampling loop
        # stores its rows into all ranks' tensors over NVLink and the kernel ends with a flag handshake.  "planes_copy": depth slabs by
        # the single-GPU kernel per view group, the slabs pushed to the peers by the copy engines under the next group's kernel
        for part in ("planes", "tiles", "planes_copy", "planes_multicast", "tiles_multicast"):
            gather = parallel.SlabGather(blk, V, 1, cfg["S"], cfg["h"], cfg["w"], dev, partition=part.split("_")[0],
                                         transport=part.split("_")[1] if "_" in part else "stores")
            if gather.mode == "all_gather" and "_" in part:      # NVLS / symmetric memory not available: variant not measured
                rec[part] = {"ms_total": None, "speedup": 0.0, "bit_exact": True, "mode": "unavailable: " + getattr(gather, "why", "")[:120]}
                continue
            corrs, masks, _ = gather(**kw)
            torch.cuda.synchronize()
            ok = all(torch.equal(a, b) for a, b in zip(corrs, full_c)) and all(torch.equal(a, b) for a, b in zip(masks, full_m))
            okt = torch.tensor([int(ok)], device=dev)
            dist.all_reduce(okt, op=dist.ReduceOp.MIN)
            ms_total = timed(lambda: gather(**kw))
            rec[part] = {"ms_total": ms_total, "speedup": ms_full / ms_total, "bit_exact": bool(okt.item()), "mode": gather.mode}
            del gather
        variants = ("planes", "tiles", "planes_copy", "planes_multicast", "tiles_multicast")
        best = max(variants, key=lambda k: rec[k]["speedup"])
        rec.update(ms_slab_total=rec[best]["ms_total"], speedup=rec[best]["speedup"], bit_exact=all(rec[k]["bit_exact"] for k in variants),
                   mode=rec[best]["mode"], partition=best)
        recs.append(rec)
    # regression with D partitioned (configs[2] cost shape)
    c3 = syn.CFG3
    m = syn.mvs_inputs(c3["h"], c3["w"], 4, 2, c3["D"], seed=5)
    D = c3["D"]
    lo, hi = parallel.slab_range(D, world, rank)
    cost, depth = cu(m["cost"]), cu(m["depth"])
    cs = cost[:, lo:hi].contiguous()
    reg = parallel.SlabRegress(1, D, c3["h"], c3["w"], dev)
    r = reg(cs, depth)
    r0 = bu.mvs_regress(cost, depth)
    rel = float(((r["depth"] - r0["depth"]).abs() / r0["depth"].abs()).max().item())
    agree = float((r["depth_index"] == r0["depth_index"]).float().mean().item())
    ms_full = timed(lambda: bu.mvs_regress(cost, depth))
    ms_slab = timed(lambda: reg(cs, depth))
    return {"what": "configs[3]: robust_mvd correlation at the ETH3D shape 96x144 (C=256, S=256, one frame) split over %d GPUs (by planes and by "
                    "tiles); every rank ends with the full [V,N,S,h,w] corr and mask (LearnedFusion reads planes as channels)" % world,
            "planesweep": recs,
            "regress": {"what": "softmax over D sharded (cost [1,192,224,304]): partials (m, Z, A, I) exchanged, merged, 4-bin confidence",
                        "ms_1gpu": ms_full, "ms_slab_total": ms_slab, "speedup": ms_full / ms_slab, "depth_rel_err": rel, "index_agreement": agree,
                        "mode": reg.mode}}


def run_ours(args):
    import torch
    import torch.distributed as dist
    from robustmvd_b200 import _lib
    from robustmvd_b200.blocks.planesweep_corr import PlanesweepCorrelation
    from robustmvd_b200.pipeline import HostPlanesweepPipeline

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        # torch / NCCL write a "NCCL version ..." banner to file descriptor 1 when the communicator comes up: point fd 1 at
        # stderr until the JSON line is printed (one-JSON-line contract of this script)
        sys.stdout.flush()
        saved_stdout = os.dup(1)
        os.dup2(2, 1)
        dist.init_process_group("nccl", device_id=dev)
    _lib.load()
    frames = args.frames
    inp = make_inputs(frames, 1234 + rank)
    host = dict(feat_key=torch.from_numpy(inp["feat_key"]).pin_memory(),
                feat_sources=[torch.from_numpy(f).pin_memory() for f in inp["feat_sources"]])
    Kk = torch.from_numpy(inp["intrinsics_key"]).to(dev)
    Ts = [torch.from_numpy(t).to(dev) for t in inp["transforms"]]
    d_key = host["feat_key"].to(dev)
    d_src = [f.to(dev) for f in host["feat_sources"]]
    block = PlanesweepCorrelation()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev) if frames < 4 else None

    def step_device():
        if flush is not None:
            flush.fill_(1)
        return block(feat_key=d_key, intrinsics_key=Kk, feat_sources=d_src, source_to_key_transforms=Ts, **KW)

    P = CFG["h"] * CFG["w"]
    out_host = torch.empty((2, CFG["V"], frames, CFG["S"], CFG["h"], CFG["w"]), dtype=torch.float32).pin_memory()

    chunks = [int(c) for c in str(args.e2e_chunk).split(',')]
    pipe = HostPlanesweepPipeline(dev, block, chunk_frames=chunks, compact_mask=True,      # mask over PCIe as bytes, expanded on the host
                                  mask_threads=args.e2e_mask_threads or None, mask_transport=args.e2e_mask_transport)
    pipe_f32 = HostPlanesweepPipeline(dev, block, chunk_frames=chunks, compact_mask=False)  # mask over PCIe as float32

    def step_e2e():   # pinned host features in, corr + mask back in pinned host memory; returns after the last byte landed
        pipe.run(host["feat_key"], Kk, host["feat_sources"], Ts, out_host[0], out_host[1], **KW)

    def step_e2e_stream():   # same, as a stream of batches: queued, consecutive batches overlap; finish() ends the timed region
        pipe.run(host["feat_key"], Kk, host["feat_sources"], Ts, out_host[0], out_host[1], wait=False, **KW)

    def step_e2e_stream_f32():
        pipe_f32.run(host["feat_key"], Kk, host["feat_sources"], Ts, out_host[0], out_host[1], wait=False, **KW)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup, finish=None):
        for _ in range(warmup):
            fn()
        if finish is not None:
            finish()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        if finish is not None:
            finish()          # host waits for the last byte; the current stream is ordered after it, so e1 follows all copies
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()) / steps

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ms_step = timed(step_device, args.steps, args.warmup)
    clocks = sampler.stop() if rank == 0 else None
    pipe_threads = pipe.mask_threads
    ms_e2e_blocking = timed(step_e2e, max(2, args.steps // 4), 2)
    ms_e2e = timed(step_e2e_stream, max(4, args.steps // 4), 3, finish=pipe.finish)
    ms_e2e_f32 = timed(step_e2e_stream_f32, max(4, args.steps // 4), 3, finish=pipe_f32.finish)
    # the compact transport must leave exactly the float32 mask of the plain copy in the host buffer (checked on the last step)
    mask_compact = out_host[1].clone()
    step_e2e_stream_f32()
    pipe_f32.finish()
    e2e_mask_equal = bool(torch.equal(mask_compact, out_host[1]))
    del mask_compact

    # per-call device time of the hot path alone (no flush), CUDA events on the launching stream
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    torch.cuda.synchronize()
    for a, b in ev:
        if flush is not None:
            flush.fill_(1)
        a.record()
        block(feat_key=d_key, intrinsics_key=Kk, feat_sources=d_src, source_to_key_transforms=Ts, **KW)
        b.record()
    torch.cuda.synchronize()
    ms_call = float(np.mean([a.elapsed_time(b) for a, b in ev]))
    peaks, which = measured_peaks()

    extra = {}
    if not args.no_secondary:
        del pipe, pipe_f32, out_host
        extra["sweep"] = sweep_records(torch, dist, dev, world, peaks)
        if world > 1:
            try:
                extra["slab"] = slab_records(torch, dist, dev, world, rank)
            except Exception as e:  # noqa: BLE001  (a failed slab leg must not lose the headline line)
                extra["slab"] = {"error": repr(e)[:400]}
        if rank == 0:
            for key, fn in (("secondary", lambda: secondary_records(torch, dev, peaks, not args.no_cpu_baseline)),
                            ("torch_gpu_baseline", lambda: torch_gpu_baseline(torch, dev)),
                            ("e2e_model", lambda: e2e_model_record(torch, dev))):
                try:
                    extra[key] = fn()
                except Exception as e:  # noqa: BLE001
                    extra[key] = {"error": repr(e)[:400]}

    if rank == 0:
        ab = algorithmic_bytes(frames)
        # duration of the launch group: the back-to-back timed region (GPU-bound, no gaps between calls) when the step is the
        # bare call; the per-call event pairs (which can include host gaps) only when an L2 flush sits between the steps
        ms_launch = ms_call if flush is not None else ms_step
        achieved = ab / (ms_launch * 1e-3) / 1e9
        h2d = 4 * frames * CFG["C"] * P * (1 + CFG["V"])
        d2h_f32 = 2 * 4 * frames * CFG["V"] * CFG["S"] * P
        mask_bits = 1 if args.e2e_mask_transport == "bits" else 8
        d2h = (4 * 8 + mask_bits) * frames * CFG["V"] * CFG["S"] * P // 8   # corr float32 + mask as bytes (or bits)
        line = base_line(args, frames, world)
        line.update({
            "impl": "ours", "value": voxels(frames) * world / (ms_step * 1e-3), "ms_per_step": ms_step,
            "frames_per_s": frames * world / (ms_step * 1e-3),
            "e2e": {"value": voxels(frames) * world / (ms_e2e * 1e-3), "unit": "voxels/s", "ms_per_step": ms_e2e,
                    "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "h2d_gbs_per_rank": h2d / ms_e2e / 1e6, "d2h_gbs_per_rank": d2h / ms_e2e / 1e6,
                    "ms_per_step_blocking": ms_e2e_blocking,
                    "value_blocking": voxels(frames) * world / (ms_e2e_blocking * 1e-3),
                    "mask_transport": "%s over PCIe (rmvd_mask_pack_%s) + host expansion to float32 (rmvd_host_expand_mask%s_async, "
                                      "%d worker threads) inside the timed region" % (
                                          (("one bit per element", "bits", "_bits") if args.e2e_mask_transport == "bits" else ("u8", "u8", ""))
                                          + (pipe_threads,)),
                    "host_mask_bit_identical_to_float32_copy": e2e_mask_equal,
                    "float32_mask_transport": {"ms_per_step": ms_e2e_f32, "value": voxels(frames) * world / (ms_e2e_f32 * 1e-3),
                                               "d2h_bytes_per_step": d2h_f32, "h2d_gbs_per_rank": h2d / ms_e2e_f32 / 1e6,
                                               "d2h_gbs_per_rank": d2h_f32 / ms_e2e_f32 / 1e6},
                    "api": "HostPlanesweepPipeline.run(wait=False) per step + finish(): PlanesweepCorrelation.forward per chunk "
                           "of %s frames from pinned host buffers, corr + mask (float32, the block's outputs bit for bit) back in "
                           "pinned host memory every step; H2D / kernels / D2H / host expansion of consecutive chunks AND of "
                           "consecutive steps overlap on 4 streams (double-buffered device inputs).  *_blocking: each step waits "
                           "for its last byte before the next starts (latency of one batch)" % args.e2e_chunk},
            "gpu_launches": args.steps * 3,   # (device-resident leg) key repack + source repack (all views) + one persistent plane-sweep kernel per step
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                         "frac": achieved / peaks["hbm_gbs"], "traffic": ncu_traffic(frames), "peak_source": which,
                         "kernel": "rmvd_psweep_corr_fwd = nchw_to_keytiles_kernel + nchw_to_grouped_kernel (split copies) + "
                                   "psweep_ws_kernel (persistent, tcgen05 + TMEM + bulk copies)",
                         "algorithmic_bytes_per_launch": ab, "ms_per_launch": ms_launch, "ms_per_call_event_pairs": ms_call},
            "clocks": clocks,
        })
        line.update(extra)
        if not args.no_cpu_baseline and world == 1:
            line["cpu_baseline"] = time_cpu_baseline(3, 1)
        if world > 1:
            sys.stdout.flush()
            os.dup2(saved_stdout, 1)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def ncu_traffic(frames):
    """DRAM bytes (read + write) of one call's kernels, summed over the launches of the call from the committed ncu --set full
    capture (profiles/roofline_traffic.json, written by tools/ncu_traffic.py from the .ncu-rep), or None when the run does not
    match the capture."""
    path = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if not os.path.exists(path):
        return None
    t = json.load(open(path))
    return t["dram_bytes_per_call"] if t.get("frames") == frames else None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--frames", type=int, default=16, help="frames per GPU per step")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-secondary", action="store_true", help="headline line only (no secondary / baseline / sweep / slab records)")
    ap.add_argument("--e2e-mask-transport", default="u8", choices=("bits", "u8"), help="compact mask transport of the e2e leg (bits measured slower)")
    ap.add_argument("--e2e-mask-threads", type=int, default=0, help="worker threads of the host-side mask expansion (0 = pipeline default)")
    ap.add_argument("--e2e-chunk", default="2", help="frames per chunk of the host pipeline (e2e leg); a,b,c = schedule, last repeats")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
