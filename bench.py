#!/usr/bin/env python
"""Benchmark of the plane-sweep cost-volume hot path (BASELINE.json metric: cost-volume voxels/s @384x1280).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--frames F]

Workload (configs[1]): robust_mvd plane-sweep correlation at the KITTI shape 384x1280 -> 48x160 features, C = 256,
S = 256 planes, 2 source views.  One step = one call of PlanesweepCorrelation.forward on a batch of F frames per GPU
(default 8: 189 MB of features in, 252 MB of corr + mask out, i.e. larger than the 126 MB L2, so consecutive steps
cannot be served from cache).  N > 1: one process per GPU (torchrun), each rank sweeps its own frames (weak scaling
by samples, no data-path collective); the time is the max over ranks.

Prints ONE JSON line (see the round contract): value = voxels/s with inputs resident in HBM; e2e = the same through
the block API from pinned HOST buffers with the H2D / D2H copies inside the timed region; roofline = algorithmic
bytes / device time against MEASURED_PEAKS.json; cpu_baseline = oracle/torch_cpu.py (the reference's operator sequence
on the host cores) on a bounded sample.
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


def voxels(frames):
    return frames * CFG["V"] * CFG["S"] * CFG["h"] * CFG["w"]


def algorithmic_bytes(frames):
    """features read once, corr + mask written once."""
    P = CFG["h"] * CFG["w"]
    return 4 * frames * CFG["C"] * P * (1 + CFG["V"]) + 2 * 4 * frames * CFG["V"] * CFG["S"] * P


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


def make_inputs(frames, seed):
    inp = syn.planesweep_inputs(CFG["h"], CFG["w"], CFG["C"], CFG["V"], n=frames, seed=seed)
    return inp


def cpu_reference_step(inp1):
    from oracle import torch_cpu
    return torch_cpu.planesweep_correlation(inp1["feat_key"], inp1["intrinsics_key"], inp1["feat_sources"],
                                            inp1["transforms"], **KW)


def time_cpu_baseline(steps, warmup):
    """The reference's algorithm on the host cores, one frame per step (bounded sample of the workload)."""
    import torch
    torch.set_num_threads(os.cpu_count())
    inp1 = make_inputs(1, 1234)
    for _ in range(warmup):
        cpu_reference_step(inp1)
    t0 = time.perf_counter()
    for _ in range(steps):
        cpu_reference_step(inp1)
    dt = (time.perf_counter() - t0) / steps
    return dict(value=voxels(1) / dt, unit="voxels/s", cores=torch.get_num_threads(), kind="port",
                sample=f"1 frame of the workload per step ({steps} steps, {dt * 1e3:.0f} ms/frame); oracle/torch_cpu.py = "
                       "the reference's matmul + grid_sample sequence on torch CPU", ms_per_frame=dt * 1e3)


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

    pipe = HostPlanesweepPipeline(dev, block, chunk_frames=[int(c) for c in str(args.e2e_chunk).split(',')])

    def step_e2e():   # pinned host features in, corr + mask back in pinned host memory; returns after the last byte landed
        pipe.run(host["feat_key"], Kk, host["feat_sources"], Ts, out_host[0], out_host[1], **KW)

    def step_e2e_stream():   # same, as a stream of batches: queued, consecutive batches overlap; finish() ends the timed region
        pipe.run(host["feat_key"], Kk, host["feat_sources"], Ts, out_host[0], out_host[1], wait=False, **KW)

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
    ms_e2e_blocking = timed(step_e2e, max(2, args.steps // 4), 2)
    ms_e2e = timed(step_e2e_stream, max(4, args.steps // 4), 3, finish=pipe.finish)

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

    if rank == 0:
        peaks, which = measured_peaks()
        ab = algorithmic_bytes(frames)
        # duration of the launch group: the back-to-back timed region (GPU-bound, no gaps between calls) when the step is the
        # bare call; the per-call event pairs (which can include host gaps) only when an L2 flush sits between the steps
        ms_launch = ms_call if flush is not None else ms_step
        achieved = ab / (ms_launch * 1e-3) / 1e9
        h2d = 4 * frames * CFG["C"] * P * (1 + CFG["V"])
        d2h = 2 * 4 * frames * CFG["V"] * CFG["S"] * P
        line = base_line(args, frames, world)
        line.update({
            "impl": "ours", "value": voxels(frames) * world / (ms_step * 1e-3), "ms_per_step": ms_step,
            "frames_per_s": frames * world / (ms_step * 1e-3),
            "e2e": {"value": voxels(frames) * world / (ms_e2e * 1e-3), "unit": "voxels/s", "ms_per_step": ms_e2e,
                    "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step_blocking": ms_e2e_blocking,
                    "value_blocking": voxels(frames) * world / (ms_e2e_blocking * 1e-3),
                    "api": "HostPlanesweepPipeline.run(wait=False) per step + finish(): PlanesweepCorrelation.forward per chunk "
                           "of %s frames from pinned host buffers, corr + mask copied back to pinned host memory every step; "
                           "H2D / kernels / D2H of consecutive chunks AND of consecutive steps overlap on 3 streams "
                           "(double-buffered device inputs).  *_blocking: each step waits for its last byte before the next "
                           "starts (latency of one batch)" % args.e2e_chunk},
            "gpu_launches": args.steps * 2,   # one repack launch (key + V source maps) + one plane-sweep kernel per step
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                         "frac": achieved / peaks["hbm_gbs"], "traffic": ncu_traffic(frames), "peak_source": which,
                         "kernel": "rmvd_psweep_corr_fwd = nchw_to_split_tile_kernel (key + V sources, one launch) + psweep_tc_kernel (tcgen05 + TMA)",
                         "algorithmic_bytes_per_launch": ab, "ms_per_launch": ms_launch, "ms_per_call_event_pairs": ms_call},
            "clocks": clocks,
        })
        if not args.no_cpu_baseline and world == 1:
            line["cpu_baseline"] = time_cpu_baseline(3, 1)
        if world > 1:
            sys.stdout.flush()
            os.dup2(saved_stdout, 1)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def ncu_traffic(frames):
    """DRAM bytes (read + write) of one call's kernels from the committed ncu --set full capture (profiles/roofline_traffic.json,
    written from the capture by hand; bytes are per call of 8 frames), or None when the run does not match the capture."""
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
    ap.add_argument("--frames", type=int, default=16, help="frames per GPU per step (16 x 240 tiles = 12.97 waves of 296 CTA slots)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--e2e-chunk", default="2", help="frames per chunk of the host pipeline (e2e leg); a,b,c = schedule, last repeats")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
