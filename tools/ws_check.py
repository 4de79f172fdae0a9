"""Diagnostics for the persistent plane-sweep kernel (psweep_ws.cu) on a B200: per case the number of mask / index mismatches
against the numpy oracle and the relative corr error, without stopping at the first failure; then the timing of the
16-frame cfg2 call (the bench workload) and of single frames.

    python tools/ws_check.py [--quick] [--time-only] [--kernel ws|band|direct]

The kernel is chosen by RMVD_PSWEEP_KERNEL (read once per process by the library): --kernel sets it before the first call.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--time-only", action="store_true")
    ap.add_argument("--kernel", default=None)
    ap.add_argument("--frames", type=int, default=16)
    args = ap.parse_args()
    if args.kernel:
        os.environ["RMVD_PSWEEP_KERNEL"] = args.kernel
    import torch
    import oracle
    from tools import cases
    from robustmvd_b200 import synthetic as syn
    from robustmvd_b200.blocks.planesweep_corr import PlanesweepCorrelation

    dev = "cuda:0"
    t = lambda a: None if a is None else torch.from_numpy(np.asarray(a)).to(dev)  # noqa: E731
    bad = 0

    def run(case, debug=True):
        kw = {k: (t(v) if isinstance(v, np.ndarray) else v) for k, v in case["kwargs"].items()}
        blk = PlanesweepCorrelation(normalize=case.get("normalize", "dim"))
        corrs, masks, inv = blk(
            feat_key=t(case["feat_key"]), intrinsics_key=t(case["intrinsics_key"]),
            feat_sources=[t(f) for f in case["feat_sources"]], source_to_key_transforms=[t(x) for x in case["transforms"]],
            intrinsics_sources=None if case["intrinsics_sources"] is None else [t(x) for x in case["intrinsics_sources"]],
            debug=debug, **kw)
        torch.cuda.synchronize()
        dbg = blk.debug
        return ([c.cpu().numpy() for c in corrs], [m.cpu().numpy() for m in masks],
                None if dbg is None else {k: v.cpu().numpy() for k, v in dbg.items()})

    if not args.time_only:
        todo = []
        for c in cases.planesweep_cases(include_large=not args.quick):
            if not c.get("warp_only", False):
                todo.append(c)
        for (h, w, hs, ws, C) in [(7, 33, 7, 33, 32), (10, 130, 10, 130, 64), (100, 8, 100, 8, 32), (70, 1, 70, 1, 32),
                                  (31, 31, 17, 45, 32), (9, 65, 20, 130, 32), (5, 64, 5, 64, 96)]:
            inp = syn.planesweep_inputs(h, w, C, 2, n=2, seed=500 + h + w, hs=hs, ws=ws)
            todo.append(dict(name="shape_%dx%d_src%dx%d_c%d" % (h, w, hs, ws, C), feat_key=inp["feat_key"],
                             intrinsics_key=inp["intrinsics_key"], feat_sources=inp["feat_sources"], transforms=inp["transforms"],
                             intrinsics_sources=None, kwargs=dict(num_sampling_points=40, min_depth=0.4, max_depth=1000.0)))
        if not args.quick:
            cfg = syn.CFG2
            inp = syn.planesweep_inputs(cfg["h"], cfg["w"], cfg["C"], cfg["V"])
            todo.append(dict(name="cfg2", feat_key=inp["feat_key"], intrinsics_key=inp["intrinsics_key"],
                             feat_sources=inp["feat_sources"], transforms=inp["transforms"], intrinsics_sources=None,
                             kwargs=dict(num_sampling_points=cfg["S"], min_depth=cfg["min_depth"], max_depth=cfg["max_depth"])))
        for case in todo:
            t0 = time.time()
            try:
                ocorrs, omasks, _, odbg = oracle.planesweep_correlation(
                    case["feat_key"], case["intrinsics_key"], case["feat_sources"], case["transforms"], case["intrinsics_sources"],
                    normalize=case.get("normalize", "dim"), return_debug=True, **case["kwargs"])
                for debug in (True, False):
                    corrs, masks, dbg = run(case, debug)
                    rec = dict(case=case["name"], debug=debug)
                    ok = True
                    for v in range(len(corrs)):
                        mm = int((masks[v] != omasks[v]).sum())
                        scale = max(float(np.abs(ocorrs[v]).max()), 1e-30)
                        d = np.abs(corrs[v] - ocorrs[v])
                        err = float(np.nanmax(d)) / scale if np.isfinite(d).any() else float("nan")
                        nbad = int((~(d <= 1e-4 * scale)).sum())
                        rec["v%d" % v] = dict(mask_mismatch=mm, corr_err=err, corr_bad=nbad, n=int(d.size))
                        if dbg is not None:
                            rec["v%d" % v]["ix0_mismatch"] = int((dbg["ix0"][v] != odbg[v]["ix0"]).sum())
                            rec["v%d" % v]["iy0_mismatch"] = int((dbg["iy0"][v] != odbg[v]["iy0"]).sum())
                            ok = ok and rec["v%d" % v]["ix0_mismatch"] == 0 and rec["v%d" % v]["iy0_mismatch"] == 0
                        ok = ok and mm == 0 and nbad == 0
                        if nbad and nbad < d.size:
                            idx = np.argwhere(~(d <= 1e-4 * scale))
                            rec["v%d" % v]["first_bad"] = idx[0].tolist()
                            rec["v%d" % v]["last_bad"] = idx[-1].tolist()
                            rec["v%d" % v]["bad_planes"] = sorted(set(idx[:, 1].tolist()))[:12]
                    rec["ok"] = ok
                    rec["sec"] = round(time.time() - t0, 2)
                    bad += 0 if ok else 1
                    print(json.dumps(rec), flush=True)
            except Exception as e:  # noqa: BLE001
                bad += 1
                print(json.dumps(dict(case=case["name"], error=repr(e)[:300])), flush=True)

    # timing: cfg2, N frames per call (the bench workload) and one frame
    cfg = syn.CFG2
    for frames in (args.frames, 1):
        inp = syn.planesweep_inputs(cfg["h"], cfg["w"], cfg["C"], cfg["V"], n=frames)
        kw = dict(feat_key=t(inp["feat_key"]), intrinsics_key=t(inp["intrinsics_key"]),
                  feat_sources=[t(f) for f in inp["feat_sources"]], source_to_key_transforms=[t(x) for x in inp["transforms"]],
                  num_sampling_points=cfg["S"], min_depth=cfg["min_depth"], max_depth=cfg["max_depth"])
        blk = PlanesweepCorrelation()
        for _ in range(5):
            blk(**kw)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 30
        a.record()
        for _ in range(reps):
            blk(**kw)
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / reps
        # device time without the host: the same call captured in a CUDA graph and replayed back to back
        ms_graph = None
        try:
            g = torch.cuda.CUDAGraph()
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                blk(**kw)
            torch.cuda.current_stream().wait_stream(side)
            with torch.cuda.graph(g):
                blk(**kw)
            for _ in range(3):
                g.replay()
            torch.cuda.synchronize()
            a.record()
            for _ in range(reps):
                g.replay()
            b.record()
            torch.cuda.synchronize()
            ms_graph = a.elapsed_time(b) / reps
        except Exception as e:  # noqa: BLE001
            ms_graph = repr(e)[:200]
        P = cfg["h"] * cfg["w"]
        nbytes = 4 * frames * cfg["C"] * P * (1 + cfg["V"]) + 8 * frames * cfg["V"] * cfg["S"] * P
        print(json.dumps(dict(timing="cfg2", kernel=os.environ.get("RMVD_PSWEEP_KERNEL", "default"), frames=frames, ms=round(ms, 4), ms_graph=ms_graph if not isinstance(ms_graph, float) else round(ms_graph, 4),
                              gbs=round(nbytes / ms / 1e6, 1), frac_of_6557=round(nbytes / ms / 1e6 / 6557.4, 4))), flush=True)
    print("ws_check bad cases:", bad)
    return 0 if bad == 0 else 1


if __name__ == "__main__":
    sys.exit(main())
