"""Per-kernel SASS opcode counts of the in-tree library (cuobjdump -sass): which kernels carry tcgen05 / TMEM / bulk-copy /
mbarrier instructions.      python tools/sass_summary.py > profiles/<round>_sass_opcodes.md"""
import collections
import os
import re
import subprocess

LIB = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "robustmvd_b200", "csrc", "librmvd_b200.so")
# UTCHMMA = tcgen05.mma, UTCBAR = tcgen05.commit, UTCCP = tcgen05.cp, UBLKCP = cp.async.bulk, UTMALDG = cp.async.bulk.tensor,
# LDTM / STTM = tcgen05.ld / st, SYNCS = mbarrier ops, LDGSTS = cp.async, USETMAXREG = setmaxnreg
OPS = ["UTCHMMA", "UTCBAR", "UTCCP", "UBLKCP", "UTMALDG", "LDTM", "STTM", "SYNCS", "ELECT", "LDGSTS", "USETMAXREG", "FFMA2", "LDS.128"]


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    cur, cnt, total = None, collections.defaultdict(collections.Counter), collections.Counter()
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            continue
        m = re.search(r"\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line) if cur and "/*" in line else None
        if m:
            total[cur] += 1
            for o in OPS:
                if m.group(1).startswith(o):
                    cnt[cur][o] += 1
    dem = lambda n: subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip().split("(")[0]  # noqa: E731
    print("SASS opcode counts per kernel of `librmvd_b200.so` (sm_100a), `cuobjdump -sass`:\n")
    print("| kernel | instructions | " + " | ".join(OPS) + " |\n|---|---|" + "---|" * len(OPS))
    for k in sorted(total, key=lambda k: -total[k]):
        print(f"| `{dem(k)}` | {total[k]} | " + " | ".join(str(cnt[k][o]) for o in OPS) + " |")


if __name__ == "__main__":
    main()
