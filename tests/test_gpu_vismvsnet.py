"""GPU parity of the Vis-MVSNet cost volume: fused homography warping + group-wise correlation
(``rmvd_homography_groupcorr_fwd``) and ``homography_warping`` alone (``rmvd_homography_warp_fwd``) against the numpy oracle
and the golden vectors frozen from the unmodified reference functions (tools/make_vis_golden.py).  Bar: 1e-4 of max|ref|."""
import numpy as np
import pytest
import torch

import oracle
from tools import cases
from golden_util import load_golden

pytestmark = pytest.mark.gpu
VIS = cases.vis_cases()
TOL = 1e-4
DEV = "cuda:0"


def _t(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(DEV)


def _rel(a, b, scale=None):
    return float(np.abs(a - b).max()) / max(float(np.abs(b).max() if scale is None else scale), 1e-30)


@pytest.mark.parametrize("case", VIS, ids=[c["name"] for c in VIS])
def test_groupwise_cost_volume_matches_oracle_and_reference(cuda_lib, case):
    from robustmvd_b200.blocks import utils as bu
    g = load_golden("vis", case["name"])
    Hs = g["Hs"]
    cost = bu.groupwise_cost_volume(_t(case["ref"]), _t(case["src"]), _t(Hs), case["groups"]).cpu().numpy()
    warped = bu.homography_warp_volume(_t(case["src"]), _t(Hs)).cpu().numpy()
    oc = oracle.groupwise_cost_volume(case["ref"], case["src"], Hs, case["groups"])
    ow = oracle.homography_warp_volume(case["src"], Hs)
    assert cost.shape == oc.shape and warped.shape == ow.shape
    assert _rel(cost, oc) <= TOL and _rel(warped, ow) <= TOL
    assert _rel(cost.reshape(-1)[::int(g["cost_stride"])], g["cost_sub"], float(g["cost_absmax"])) <= TOL
    assert _rel(warped.reshape(-1)[::int(g["warped_stride"])], g["warped_sub"], float(g["warped_absmax"])) <= TOL


@pytest.mark.parametrize("case", VIS, ids=[c["name"] for c in VIS])
def test_get_homographies_mirror_matches_reference(cuda_lib, case):
    """Host-side mirror of get_homographies (blocks/utils.py:96-150), evaluated by torch on the device (3x3 products and one
    3x3 inverse per sample): within 1e-4 of the reference's matrices, and the cost volume built from it within 1e-3."""
    from robustmvd_b200.blocks import utils as bu
    g = load_golden("vis", case["name"])
    Hs = bu.get_homographies(_t(case["left_cam"]), _t(case["right_cam"]), case["depth_num"], _t(case["depth_start"]),
                             _t(case["depth_interval"]), inv=case["inv"])
    assert tuple(Hs.shape) == g["Hs"].shape
    assert _rel(Hs.cpu().numpy(), g["Hs"]) <= 1e-4
    cost = bu.groupwise_cost_volume(_t(case["ref"]), _t(case["src"]), Hs, case["groups"]).cpu().numpy()
    assert _rel(cost.reshape(-1)[::int(g["cost_stride"])], g["cost_sub"], float(g["cost_absmax"])) <= 1e-3


def test_single_homography_warping(cuda_lib):
    """homography_warping(input [n,c,h,w], H [n,3,3]) and the per-pixel form H [n,h,w,3,3] (blocks/utils.py:179-189)."""
    from robustmvd_b200.blocks import utils as bu
    case = VIS[2]   # per-pixel homographies
    Hs = load_golden("vis", case["name"])["Hs"]
    ow = oracle.homography_warp_volume(case["src"], Hs)
    got = bu.homography_warping(_t(case["src"]), _t(Hs[:, 1])).cpu().numpy()            # n h w 3 3
    assert _rel(got, ow[:, :, 1]) <= TOL
    case = VIS[0]
    Hs = load_golden("vis", case["name"])["Hs"]
    ow = oracle.homography_warp_volume(case["src"], Hs)
    got = bu.homography_warping(_t(case["src"]), _t(Hs[:, 3, 0, 0])).cpu().numpy()      # n 3 3
    assert _rel(got, ow[:, :, 3]) <= TOL


def test_identity_homography_reproduces_the_source(cuda_lib):
    """Size-independent property at a DTU-sized map: H = I samples every pixel centre, so warped == src and
    cost[g] == sum over the group of ref * src."""
    from robustmvd_b200.blocks import utils as bu
    n, c, h, w, G = 1, 32, 128, 160, 8
    gen = torch.Generator(device=DEV).manual_seed(3)
    ref = torch.randn((n, c, h, w), device=DEV, generator=gen)
    src = torch.randn((n, c, h, w), device=DEV, generator=gen)
    Hs = torch.eye(3, device=DEV).view(1, 1, 1, 1, 3, 3).repeat(n, 4, 1, 1, 1, 1)
    warped = bu.homography_warp_volume(src, Hs)
    assert float((warped - src[:, :, None]).abs().max()) <= 1e-5 * float(src.abs().max())
    cost = bu.groupwise_cost_volume(ref, src, Hs, G)
    want = (ref * src).view(n, G, c // G, h, w).sum(2)
    assert float((cost - want[:, :, None]).abs().max()) <= 1e-5 * float(want.abs().max())


def test_shape_validation(cuda_lib):
    import ctypes
    one = ctypes.c_void_p(256)
    st = (ctypes.c_int64 * 4)(9, 9, 0, 0)
    assert cuda_lib.rmvd_homography_groupcorr_fwd(one, one, one, st, 1, 10, 4, 4, 2, 4, one, None) == -2   # C % groups != 0
    assert cuda_lib.rmvd_homography_warp_fwd(None, one, st, 1, 8, 4, 4, 2, one, None) == -1


@pytest.mark.parametrize("case", VIS, ids=[c["name"] for c in VIS])
def test_cost_volume_straight_from_the_cameras(cuda_lib, case):
    """``rmvd_vis_cost_volume_fwd``: get_homographies folded into the kernel (H = A - B / (depth + 1e-9), A and B per sample in
    double).  Against the reference's cost volume, whose own homographies carry float32 rounding of a longer chain: 5e-4 of
    max|ref| (the ABI path that takes the reference's Hs tensor holds 1e-4, above)."""
    from robustmvd_b200.blocks import utils as bu
    g = load_golden("vis", case["name"])
    cost = bu.vis_cost_volume(_t(case["ref"]), _t(case["src"]), _t(case["left_cam"]), _t(case["right_cam"]), case["depth_num"],
                              _t(case["depth_start"]), _t(case["depth_interval"]), case["groups"], case["inv"]).cpu().numpy()
    assert _rel(cost.reshape(-1)[::int(g["cost_stride"])], g["cost_sub"], float(g["cost_absmax"])) <= 5e-4
