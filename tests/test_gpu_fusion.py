"""GPU parity of the learned view fusion kernel (csrc/fusion.cu, through the C ABI via ``LearnedFusion``) against the numpy
oracle and the golden vectors frozen from the unmodified reference module (tools/make_fusion_golden.py).  Fused mask
bit-exact; view weights and fused correlation within 1e-5 of max|ref| (float32 softmax / division)."""
import glob
import os

import numpy as np
import pytest
import torch

import oracle
from golden_util import GOLDEN

pytestmark = pytest.mark.gpu
FILES = sorted(glob.glob(os.path.join(GOLDEN, "fusion_*.npz")))
TOL = 1e-5


def _fuse_cuda(logits, corrs, masks):
    """Run LearnedFusion's CUDA path on given logits: the module's convolutions are replaced by a lookup of the logits."""
    from robustmvd_b200.models.nets import LearnedFusion
    dev = "cuda:0"
    blk = LearnedFusion().to(dev)
    tl = [torch.from_numpy(np.ascontiguousarray(x)).to(dev) for x in logits]
    tc = [torch.from_numpy(np.ascontiguousarray(x)).to(dev) for x in corrs]
    tm = [torch.from_numpy(np.ascontiguousarray(x)).to(dev) for x in masks]
    by_ptr = {c.data_ptr(): l for c, l in zip(tc, tl)}

    class Lookup(torch.nn.Module):
        def forward(self, c):
            return by_ptr[c.data_ptr()]

    blk.corr_to_view_weight = Lookup()
    fc, fm = blk(tc, tm)
    torch.cuda.synchronize()
    w = None if len(corrs) == 1 else np.stack([x.cpu().numpy() for x in blk.view_weights], 0)
    return w, fc.cpu().numpy(), fm.cpu().numpy()


def _check(w, fc, fm, ow, ofc, ofm):
    assert np.array_equal(fm, ofm)
    scale = max(float(np.abs(ofc).max()), 1e-30)
    assert float(np.abs(fc - ofc).max()) <= TOL * scale
    if ow is not None:
        assert float(np.abs(w - ow).max()) <= TOL


@pytest.mark.parametrize("path", FILES, ids=[os.path.basename(f)[7:-4] for f in FILES])
def test_fusion_matches_oracle_and_golden(cuda_lib, path):
    g = np.load(path)
    V = int(g["V"])
    logits = [g[f"logit_{v}"] for v in range(V)]
    corrs = [g[f"corr_{v}"] for v in range(V)]
    masks = [g[f"mask_{v}"].astype(np.float32) for v in range(V)]
    w, fc, fm = _fuse_cuda(logits, corrs, masks)
    ow, ofc, ofm = oracle.view_fusion(logits, corrs, masks)
    _check(w, fc, fm, ow, ofc, ofm)
    gw = None if V == 1 else np.stack([g[f"weight_{v}"] for v in range(V)], 0)
    _check(w, fc, fm, gw, g["fused_corr"], g["fused_mask"])


@pytest.mark.parametrize("shape", [(2, 2, 256, 48, 160), (10, 1, 256, 24, 36), (3, 1, 13, 7, 33), (64, 1, 5, 3, 9)],
                         ids=["cfg2_2frames", "v10", "ragged_planes", "v64"])
def test_fusion_sizes_against_oracle(cuda_lib, shape):
    """Bench shape, many views, a plane count that is not a multiple of the planes per thread, the view limit."""
    V, N, S, h, w = shape
    rng = np.random.default_rng(V * 1000 + S)
    logits = [rng.standard_normal((N, 1, h, w)).astype(np.float32) * 4 for _ in range(V)]
    masks = [(rng.random((N, S, h, w)) < 0.5).astype(np.float32) for _ in range(V)]
    corrs = [rng.standard_normal((N, S, h, w)).astype(np.float32) * m for m in masks]
    wv, fc, fm = _fuse_cuda(logits, corrs, masks)
    _check(wv, fc, fm, *oracle.view_fusion(logits, corrs, masks))
    # property: the fused mask is the union of the view masks, and fused corr vanishes outside it
    union = (np.stack(masks, 0).sum(0) > 0).astype(np.float32)
    assert np.array_equal(fm, union) and not np.any(fc[union == 0])


def test_fusion_rejects_too_many_views(cuda_lib):
    from robustmvd_b200 import _lib
    x = torch.zeros((1, 1, 2, 2), device="cuda:0")
    c = torch.zeros((1, 3, 2, 2), device="cuda:0")
    out = torch.empty_like(c)
    st = _lib.load().rmvd_view_fusion_fwd(_lib.ptr_array([x] * 65), _lib.ptr_array([c] * 65), _lib.ptr_array([c] * 65), 65, 1, 3,
                                          2, 2, None, _lib.ptr(out), _lib.ptr(out), _lib.stream_ptr())
    assert st < 0
