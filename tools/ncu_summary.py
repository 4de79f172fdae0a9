"""Markdown summary of an ncu report (``ncu --set full``): the metrics the round's roofline discussion uses.
    python tools/ncu_summary.py gpurun_out/<x>.ncu-rep [more.ncu-rep ...] > profiles/<name>.md
"""
import csv
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smsp__cycles_active.avg",
]


def summarise(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    for vals in rows[2:]:
        d = {h: (v, u) for h, u, v in zip(hdr, units, vals)}
        print(f"### `{d['Kernel Name'][0]}`  ({rep.split('/')[-1]})\n")
        print("| metric | value | unit |\n|---|---|---|")
        for k in KEYS:
            if k in d and d[k][0] != "":
                print(f"| `{k}` | {d[k][0]} | {d[k][1]} |")
        stalls = sorted(((float(v[0]), h) for h, v in d.items()
                         if h.startswith("smsp__pcsamp_warps_issue_stalled") and "not_issued" not in h and v[0]), reverse=True)
        tot = sum(s for s, _ in stalls) or 1.0
        print("\nWarp-state samples (top): " + ", ".join(f"{h.replace('smsp__pcsamp_warps_issue_stalled_', '')} {100 * s / tot:.0f} %"
                                                          for s, h in stalls[:8]) + "\n")


if __name__ == "__main__":
    for r in sys.argv[1:]:
        summarise(r)
