"""rmvd_psweep_corr_fwd_static (the allocation-free variant: scratch in a static device array inside the library) against
rmvd_psweep_corr_fwd (scratch in a caller's workspace), both through the C ABI, on calls in a row that share the array.
Bars as in test_gpu_planesweep.py: masks bit-exact, corr within 1e-4 of max|corr|.  Collected last on purpose (the file
name)."""
import numpy as np
import pytest
import torch

from robustmvd_b200 import synthetic as syn

pytestmark = pytest.mark.gpu
TOL = 1e-4
STATIC_LIMIT = 32 << 20      # RMVD_STATIC_SCRATCH_BYTES of include/rmvd_b200.h
E_STATIC_LIMIT = -6


def _args(h, w, C, V, N, S, seed):
    from robustmvd_b200 import _lib
    dev = "cuda:0"
    inp = syn.planesweep_inputs(h, w, C, V, n=N, seed=seed)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).to(dev)  # noqa: E731
    key, Kk = t(inp["feat_key"]), t(inp["intrinsics_key"])
    srcs = [t(f) for f in inp["feat_sources"]]
    Ts = [t(x) for x in inp["transforms"]]
    Ks = [Kk] * V
    inv = t(np.linspace(1.0 / 1000.0, 1.0 / 0.4, S))
    hold = dict(key=key, Kk=Kk, srcs=srcs, Ts=Ts, inv=inv)      # keeps the device tensors alive
    common = [_lib.ptr(key), _lib.ptr_array(srcs), _lib.ptr(Kk), _lib.ptr_array(Ks), _lib.ptr_array(Ts), _lib.ptr(inv),
              _lib.i64_array([0, 1, 0, 0]), N, C, h, w, V, _lib.int_array([h] * V), _lib.int_array([w] * V), S, 0, S, 1]
    return hold, common, (V, N, S, h, w)


def _existing(lib, common, shape):
    from robustmvd_b200 import _lib
    V, N, S, h, w = shape
    C = common[8]
    corr = torch.full(shape, float("nan"), device="cuda:0")
    mask = torch.full(shape, float("nan"), device="cuda:0")
    nb = lib.rmvd_psweep_workspace_bytes(N, C, h, w, V, common[12], common[13])
    wsp = torch.empty(nb, dtype=torch.uint8, device="cuda:0")
    st = lib.rmvd_psweep_corr_fwd(*common, _lib.ptr(corr), _lib.ptr(mask), S, 0, None, None, None, _lib.ptr(wsp), nb,
                                  _lib.stream_ptr())
    _lib.check(st, "rmvd_psweep_corr_fwd")
    torch.cuda.synchronize()
    return corr.cpu().numpy(), mask.cpu().numpy(), nb


def _static(lib, common, shape):
    from robustmvd_b200 import _lib
    S = shape[2]
    corr = torch.full(shape, float("nan"), device="cuda:0")
    mask = torch.full(shape, float("nan"), device="cuda:0")
    st = lib.rmvd_psweep_corr_fwd_static(*common, _lib.ptr(corr), _lib.ptr(mask), S, 0, _lib.stream_ptr())
    torch.cuda.synchronize()
    return st, corr.cpu().numpy(), mask.cpu().numpy()


def _same(a, b):
    corr_a, mask_a = a
    corr_b, mask_b = b
    assert np.isfinite(corr_a).all() and np.isfinite(mask_a).all(), "the static call left outputs unwritten"
    assert np.array_equal(mask_a, mask_b), f"{(mask_a != mask_b).sum()} mask mismatches"
    scale = max(float(np.abs(corr_b).max()), 1e-30)
    err = float(np.abs(corr_a - corr_b).max()) / scale
    assert err <= TOL, f"corr differs by {err} of max|corr|"


# (h, w, C, V, N, S): persistent tcgen05 kernel on a small map, the KITTI shape of BASELINE.json configs[1] at one frame
# (23.6 MB of the 32 MiB), a ragged shape, and a channel count that takes the general direct kernel
SHAPES = [(24, 80, 64, 2, 2, 48), (48, 160, 256, 2, 1, 256), (17, 45, 32, 3, 2, 40), (12, 40, 24, 2, 1, 16)]


def test_static_scratch_matches_workspace_call_on_calls_in_a_row(cuda_lib):
    lib = cuda_lib
    expected = []
    for k, (h, w, C, V, N, S) in enumerate(SHAPES):
        hold, common, shape = _args(h, w, C, V, N, S, seed=900 + k)
        corr, mask, nb = _existing(lib, common, shape)
        assert nb <= STATIC_LIMIT
        expected.append((hold, common, shape, corr, mask))
    # every shape twice, in two different orders: each call finds the previous call's contents in the shared array
    order = list(range(len(SHAPES))) + list(reversed(range(len(SHAPES)))) + [1, 1]
    for k in order:
        hold, common, shape, corr, mask = expected[k]
        st, c2, m2 = _static(lib, common, shape)
        assert st == 0, lib.rmvd_status_string(st)
        _same((c2, m2), (corr, mask))


def test_static_scratch_rejects_sizes_beyond_the_limit(cuda_lib):
    lib = cuda_lib
    h, w, C, V, N, S = 48, 160, 256, 2, 2, 8         # two frames of the KITTI shape need 47 MB
    hold, common, shape = _args(h, w, C, V, N, S, seed=77)
    assert lib.rmvd_psweep_workspace_bytes(N, C, h, w, V, common[12], common[13]) > STATIC_LIMIT
    st, corr, mask = _static(lib, common, shape)
    assert st == E_STATIC_LIMIT
    assert b"static scratch" in lib.rmvd_status_string(st)
    assert np.isnan(corr).all() and np.isnan(mask).all()      # nothing was launched
    # and the array still serves the next call that fits
    hold, common, shape = _args(24, 80, 64, 2, 2, 48, seed=900)
    corr, mask, _ = _existing(lib, common, shape)
    st, c2, m2 = _static(lib, common, shape)
    assert st == 0
    _same((c2, m2), (corr, mask))


def test_static_call_allocates_no_device_memory(cuda_lib):
    """cudaMemGetInfo around the call: the free memory of the device does not move (the scratch is part of the module image)."""
    lib = cuda_lib
    hold, common, shape = _args(24, 80, 64, 2, 2, 48, seed=900)
    st, _, _ = _static(lib, common, shape)              # set-up call: symbol address, kernel attributes
    assert st == 0
    from robustmvd_b200 import _lib
    corr = torch.empty(shape, device="cuda:0")
    mask = torch.empty(shape, device="cuda:0")
    torch.cuda.synchronize()
    before = torch.cuda.mem_get_info()[0]
    stats0 = torch.cuda.memory_stats().get("allocation.all.allocated", 0)
    for _ in range(3):
        st = lib.rmvd_psweep_corr_fwd_static(*common, _lib.ptr(corr), _lib.ptr(mask), shape[2], 0, _lib.stream_ptr())
        assert st == 0
    torch.cuda.synchronize()
    assert torch.cuda.memory_stats().get("allocation.all.allocated", 0) == stats0
    assert torch.cuda.mem_get_info()[0] == before
