"""Builds the sm_100a C-ABI library ``csrc/librmvd_b200.so`` in-tree with nvcc (no torch types involved)."""
import os
import shutil
import subprocess

CSRC = os.path.join(os.path.dirname(os.path.abspath(__file__)), "csrc")
LIB = os.path.join(CSRC, "librmvd_b200.so")
SOURCES = ("psweep.cu", "psweep_band.cu", "psweep_ws.cu", "psweep_push.cu", "psweep_bwd.cu", "mvs.cu", "mvs_var.cu", "mvs_bwd.cu", "fusion.cu", "vis.cu", "hostpipe.cu")
HEADERS = ("common.cuh", "psweep.cuh", "mvs_coord.cuh", os.path.join("..", "..", "include", "rmvd_b200.h"))
NVCC_FLAGS = ["-O3", "-std=c++17", "--threads", "0", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
              "-Xcompiler", "-fPIC", "-Xcompiler", "-pthread", "-shared"]


def _nvcc():
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(exe):
        raise RuntimeError("nvcc not found: the CUDA library cannot be built")
    return exe


def is_stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, s) for s in SOURCES + HEADERS]
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def build(force=False, verbose=False):
    """Compile the library if missing or older than its sources.  Returns the path of the .so."""
    if not force and not is_stale():
        return LIB
    defs = ["-D" + d for d in os.environ.get("RMVD_BUILD_DEFS", "").split() if d]   # e.g. RMVD_WS_WATCHDOG (development builds)
    cmd = [_nvcc()] + NVCC_FLAGS + defs + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB] + list(SOURCES)
    res = subprocess.run(cmd, cwd=CSRC, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
    if verbose:
        print(res.stderr)
    return LIB


if __name__ == "__main__":
    print(build(force=True, verbose=True))
