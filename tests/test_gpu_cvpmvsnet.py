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
