"""Per-source-line summary of an ncu report captured with --import-source on (kernels compiled with -lineinfo):
    python tools/ncu_lines.py gpurun_out/<x>.ncu-rep [top]
Prints, for the hottest source lines, the share of executed warp instructions and of stall samples."""
import csv
import subprocess
import sys


def main():
    rep = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    fname, lines = "?", []
    for r in rows:
        if len(r) >= 2 and r[0] == "File Path":
            fname = r[1].split("/")[-1]
        elif len(r) > 8 and r[0].isdigit():
            try:
                lines.append((fname, int(r[0]), r[1].strip(), float(r[6] or 0), float(r[7] or 0), r))
            except ValueError:
                pass
    tot_i = sum(x[4] for x in lines) or 1
    tot_s = sum(x[3] for x in lines) or 1
    print(f"total warp instructions {tot_i:.3e}, stall samples {tot_s:.0f}")
    for f, ln, src, smp, ins, r in sorted(lines, key=lambda x: -x[3])[:top]:
        print(f"{f}:{ln:<4} inst {ins / tot_i * 100:5.1f}%  samples {smp / tot_s * 100:5.1f}%  {src[:100]}")


if __name__ == "__main__":
    main()
