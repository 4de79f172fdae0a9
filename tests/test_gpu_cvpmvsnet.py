"""GPU parity of CVP-MVSNet's cost-volume functions: ``proj_cost`` (``rmvd_cvp_proj_cost_fwd``, per-pixel
depth hypotheses, reference in-place quirk included) and ``homo_warping`` against the numpy oracle and golden vectors frozen
from the unmodified reference functions (tools/make_cvp_golden.py).  Bar: 1e-4 of max|ref|."""
import types

import numpy as np
import pytest
import torch

import oracle
from tools import cases
from golden_util import load_golden

pytestmark = pytest.mark.gpu
CVP = cases.cvp_cases()
TOL = 1e-4
DEV = "cuda:0"


def _t(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(DEV)


def _rel(a, b):
    return float(np.abs(a - b).max()) / max(float(np.abs(b).max()), 1e-30)


@pytest.mark.parametrize("c", CVP, ids=[c["name"] for c in CVP])
def test_proj_cost_and_homo_warping_match_oracle_and_reference(cuda_lib, c):
    from robustmvd_b200.blocks import cvp_mvsnet_components as cvp
    g = load_golden("cvp", c["name"])
    settings = types.SimpleNamespace(nsrc=c["nsrc"], mode="test")
    ref = _t(c["ref"])
    ref_before = ref.clone()
    cost = cvp.proj_cost(settings, ref, [[_t(f)] for f in c["srcs"]], 0, _t(c["ref_in"]), _t(c["src_in"]), _t(c["ref_ex"]),
                         _t(c["src_ex"]), _t(c["depth_hypos"])).cpu().numpy()
    assert torch.equal(ref, ref_before)            # inputs are never written
    oc = oracle.cvpmvsnet.proj_cost(c["ref"], c["srcs"], c["ref_in"], c["src_in"], c["ref_ex"], c["src_ex"], c["depth_hypos"])
    assert cost.shape == g["cost"].shape
    assert _rel(cost, oc) <= TOL and _rel(cost, g["cost"]) <= TOL
    warped = cvp.homo_warping(_t(c["srcs"][0]), _t(c["ref_in"]), _t(c["src_in"][:, 0]), _t(c["ref_ex"]), _t(c["src_ex"][:, 0]),
                              _t(c["depth_planes"])).cpu().numpy()
    assert warped.shape == g["warped"].shape and _rel(warped, g["warped"]) <= TOL


def test_proj_cost_quirk_is_the_references(cuda_lib):
    """With identity geometry (source = reference camera) every warped map equals its source map at interior pixels, so
    cost = (r^2 + sum s_v^2)/V - ((r^2 + sum s_v)/V)^2 -- the reference's in-place ``pow_`` (:392), not the variance."""
    from robustmvd_b200.blocks import cvp_mvsnet_components as cvp
    B, C, H, W, D, nsrc = 1, 8, 16, 20, 3, 2
    gen = torch.Generator(device=DEV).manual_seed(5)
    ref = torch.randn((B, C, H, W), device=DEV, generator=gen)
    srcs = [torch.randn((B, C, H, W), device=DEV, generator=gen) for _ in range(nsrc)]
    K = torch.tensor([[18.0, 0, 10.0], [0, 18.0, 8.0], [0, 0, 1.0]], device=DEV).repeat(B, 1, 1)
    E = torch.eye(4, device=DEV).repeat(B, 1, 1)
    hyp = torch.full((B, D, H, W), 0.7, device=DEV)
    cost = cvp.proj_cost(types.SimpleNamespace(nsrc=nsrc, mode="test"), ref, [[s] for s in srcs], 0, K,
                         K[:, None].repeat(1, nsrc, 1, 1), E, E[:, None].repeat(1, nsrc, 1, 1), hyp)
    V = nsrc + 1
    want = (ref ** 2 + sum(s ** 2 for s in srcs)) / V - ((ref ** 2 + sum(srcs)) / V) ** 2
    assert float((cost[:, :, 1] - want).abs().max()) <= 1e-4 * float(want.abs().max())
