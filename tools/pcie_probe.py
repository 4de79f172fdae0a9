"""PCIe rates of this box with pinned host memory: host-to-device alone, device-to-host alone, both directions at once (two
streams).  The floor of the end-to-end leg of bench.py is max(h2d_bytes / h2d_duplex, d2h_bytes / d2h_duplex).  One JSON line.

    python tools/pcie_probe.py [--mb 256] [--reps 10]
"""
import argparse
import json

import torch


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mb", type=int, default=256)
    ap.add_argument("--reps", type=int, default=10)
    args = ap.parse_args()
    n = args.mb << 20
    h_in, h_out = torch.empty(n, dtype=torch.uint8).pin_memory(), torch.empty(n, dtype=torch.uint8).pin_memory()
    d_in, d_out = torch.empty(n, dtype=torch.uint8, device="cuda"), torch.ones(n, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

    def run(h2d, d2h):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        s1.wait_stream(torch.cuda.current_stream())
        s2.wait_stream(torch.cuda.current_stream())
        for _ in range(args.reps):
            if h2d:
                with torch.cuda.stream(s1):
                    d_in.copy_(h_in, non_blocking=True)
            if d2h:
                with torch.cuda.stream(s2):
                    h_out.copy_(d_out, non_blocking=True)
        torch.cuda.current_stream().wait_stream(s1)
        torch.cuda.current_stream().wait_stream(s2)
        e1.record()
        torch.cuda.synchronize()
        return n * args.reps / (e0.elapsed_time(e1) * 1e-3) / 1e9

    for _ in range(2):
        run(True, True)
    rec = {"buffer_mb": args.mb, "h2d_alone_gbs": run(True, False), "d2h_alone_gbs": run(False, True), "each_direction_duplex_gbs": run(True, True)}
    print(json.dumps(rec))


if __name__ == "__main__":
    main()
