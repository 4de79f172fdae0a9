"""CPU: the C-ABI library builds for sm_100a, loads, and exports every symbol include/rmvd_b200.h declares.
No compute call is made here (there is no GPU); argument validation that returns before any CUDA call is."""
import ctypes
import os
import re
import shutil

import pytest

from robustmvd_b200 import _build, _lib

HEADER = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "rmvd_b200.h")


@pytest.fixture(scope="module")
def lib():
    if _build.is_stale() and (shutil.which("nvcc") or os.path.exists("/usr/local/cuda/bin/nvcc")):
        _build.build()
    return _lib.load()


def declared_functions():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(rmvd_[a-z0-9_]+)\s*\(", text)))


def test_header_and_binding_agree():
    assert declared_functions() == sorted(_lib.EXPORTS)


def test_every_declared_symbol_is_exported(lib):
    for name in declared_functions():
        assert hasattr(lib, name), name
    assert lib.rmvd_abi_version() == _lib.ABI_VERSION == int(re.search(r"#define RMVD_ABI_VERSION (\d+)", open(HEADER).read()).group(1))


def test_only_sm100a_code_is_embedded():
    exe = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(exe):
        pytest.skip("cuobjdump not available")
    import subprocess
    out = subprocess.run([exe, "-lelf", _build.LIB], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_\d+a?", out))
    assert archs == {"sm_100a"}, archs


def test_argument_validation_needs_no_device(lib):
    null = ctypes.c_void_p(None)
    st = (ctypes.c_int64 * 4)(0, 1, 0, 0)
    assert lib.rmvd_softargmax_combine(null, 2, 1, 4, 4, null, null) == -1          # RMVD_E_NULL
    one = ctypes.c_void_p(1)
    assert lib.rmvd_softargmax_combine(one, 0, 1, 4, 4, one, null) == -2            # RMVD_E_SHAPE
    assert lib.rmvd_mvs_variance_fwd(one, one, one, 1, 3, 8, 4, 4, 8, 2, 9, one, 8, 0, null) == -5  # RMVD_E_RANGE
    assert lib.rmvd_warp_grid_fwd(one, one, 1, 1, 4, 4, 4, 4, 7, 1, one, one, null) == -3  # RMVD_E_MODE
    assert lib.rmvd_mvs_variance_fwd_ws(one, one, one, 1, 3, 32, 4, 4, 8, 0, 8, one, 8, 0, null, 0, null) == -1   # RMVD_E_NULL: workspace needed
    assert lib.rmvd_mvs_variance_workspace_bytes(2, 5, 32, 7, 9) == 4 * 2 * 4 * 32 * 63
    assert lib.rmvd_volume_reduce_fwd(one, one, 1, 4, 4, 4, 3, one, null) == -3                                   # RMVD_E_MODE
    assert lib.rmvd_copy_rows(one, one, 16, 32, 1, null) == -2                                                    # row wider than the pitch
    assert lib.rmvd_sampling_points_fwd(one, one, one, one, st, 1, 4, 4, 4, 4, 4, null, null, null, null, null, null, null) == -1
    assert lib.rmvd_status_string(-4).decode().startswith("workspace")
    arr = (ctypes.c_void_p * 1)(1)
    ints = (ctypes.c_int * 1)(4)
    st = (ctypes.c_int64 * 4)(0, 1, 0, 0)
    assert lib.rmvd_psweep_corr_bwd(one, arr, one, arr, arr, one, st, 1, 8, 4, 4, 1, ints, ints, 4, 0, 4, 2, one, 4, 0,
                                    one, arr, one, 1 << 20, null) == -3             # NORM_BEFORE: caller differentiates
    assert lib.rmvd_psweep_corr_bwd(one, arr, one, arr, arr, one, st, 1, 8, 4, 4, 1, ints, ints, 4, 0, 4, 1, one, 4, 0,
                                    one, arr, one, 16, null) == -4                  # RMVD_E_WORKSPACE
    assert lib.rmvd_psweep_bwd_workspace_bytes(1, 8, 4, 4, 1, ints, ints) == 3 * 512
    assert lib.rmvd_homography_groupcorr_fwd(one, one, one, st, 1, 10, 4, 4, 2, 4, one, null) == -2  # C % groups != 0
    # push transport of the depth slabs: queue size, same stream twice, queue too small (all before any device work)
    q1, q2 = lib.rmvd_psweep_push_queue_bytes(1, 96, 144, 1, 128), lib.rmvd_psweep_push_queue_bytes(1, 96, 144, 2, 128)
    assert q1 > 16 and (q1 - 16) % 80 == 0 and q2 - 16 == 2 * (q1 - 16) and lib.rmvd_psweep_push_queue_bytes(1, 96, 144, 0, 8) == 0
    big = ctypes.c_void_p(4096)
    push = lambda q_bytes, s0, s1, ctas: lib.rmvd_psweep_corr_fwd_push(  # noqa: E731
        one, arr, one, arr, arr, one, st, 1, 64, 4, 4, 1, ints, ints, 4, 0, 4, 1, 4, 0, 1, 0, arr, arr, arr, 1, big, q_bytes, ctas, one, 1 << 20, s0, s1)
    assert push(1 << 20, null, null, 8) == -3            # RMVD_E_MODE: the pusher cannot share the sweep's stream
    assert push(64, null, one, 8) == -4                  # RMVD_E_WORKSPACE: queue smaller than rmvd_psweep_push_queue_bytes
    assert push(1 << 20, null, one, 0) == -5             # RMVD_E_RANGE: no pusher CTAs


def test_host_mask_expansion(lib):
    """rmvd_host_expand_mask is host code (the worker pool of the compact mask transport): bytes -> float32, any alignment, any
    thread count, nothing written outside [0, n)."""
    import numpy as np
    rng = np.random.default_rng(3)
    assert lib.rmvd_host_expand_mask(None, None, 4, 2) == -1
    for n in (0, 1, 15, 16, 17, 3 * 65536 + 5, 1 << 21):
        src = (rng.random(n + 3) < 0.5).astype(np.uint8)
        for off in (0, 1, 3):
            for threads in (1, 3, 8):
                dst = np.full(n + 8, 7.0, np.float32)
                s, d = src[off:off + n], dst[off:off + n]
                assert lib.rmvd_host_expand_mask(ctypes.c_void_p(s.ctypes.data), ctypes.c_void_p(d.ctypes.data), n, threads) == 0
                assert np.array_equal(d, s.astype(np.float32)) and np.all(dst[:off] == 7) and np.all(dst[off + n:] == 7)


def test_host_mask_bit_expansion(lib):
    """rmvd_host_expand_mask_bits: one bit per element (numpy packbits, little bit order) -> float32, any alignment and thread count."""
    import numpy as np
    rng = np.random.default_rng(4)
    for n in (1, 7, 8, 31, 32, 33, 3 * 65536 + 5, 1 << 21):
        m = rng.random(n) < 0.5
        packed = np.packbits(m, bitorder="little")
        pad = np.zeros((len(packed) + 3) // 4 * 4, np.uint8)
        pad[:len(packed)] = packed
        for off in (0, 1):
            for threads in (1, 3, 8):
                dst = np.full(n + 8, 7.0, np.float32)
                d = dst[off:off + n]
                assert lib.rmvd_host_expand_mask_bits(ctypes.c_void_p(pad.ctypes.data), ctypes.c_void_p(d.ctypes.data), n, threads) == 0
                assert np.array_equal(d, m.astype(np.float32)) and np.all(dst[:off] == 7) and np.all(dst[off + n:] == 7)


def test_host_mask_expansion_survives_fork(lib):
    """The worker pool belongs to the process that created it: a forked child (e.g. a data-loader worker) builds its own."""
    import numpy as np
    n = 1 << 20
    src = (np.arange(n) % 3 == 0).astype(np.uint8)
    dst = np.empty(n, np.float32)
    assert lib.rmvd_host_expand_mask(ctypes.c_void_p(src.ctypes.data), ctypes.c_void_p(dst.ctypes.data), n, 4) == 0
    pid = os.fork()
    if pid == 0:
        ok = 1
        try:
            out = np.empty(n, np.float32)
            rc = lib.rmvd_host_expand_mask(ctypes.c_void_p(src.ctypes.data), ctypes.c_void_p(out.ctypes.data), n, 4)
            ok = 0 if rc == 0 and np.array_equal(out, src.astype(np.float32)) else 1
        finally:
            os._exit(ok)
    _, status = os.waitpid(pid, 0)
    assert os.WIFEXITED(status) and os.WEXITSTATUS(status) == 0
    assert np.array_equal(dst, src.astype(np.float32))


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_build, "LIB", str(tmp_path / "nope.so"))
    with pytest.raises(_lib.RmvdError):
        _lib.load()


def test_cpu_tensors_are_rejected():
    import torch
    with pytest.raises(_lib.RmvdError):
        _lib.ptr(torch.zeros(4))
