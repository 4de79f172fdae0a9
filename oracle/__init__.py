"""CPU oracle for the plane-sweep cost-volume hot path -- TEST INFRASTRUCTURE ONLY.

This package is a numpy restatement of the reference's algorithm
(``/root/reference/rmvd/models/blocks/planesweep_corr.py``,
``/root/reference/rmvd/models/blocks/utils.py:222-274`` and
``/root/reference/rmvd/models/mvsnet.py:123-160``).  Each function cites the
reference ``file:line`` it follows.

Added for the "next" rows of SURVEY.md 8(f): ``planesweep_correlation_backward`` (what autograd returns through
``TorchCorr.forward``, pinned by ``tools/make_bwd_golden.py``), ``vismvsnet`` (``blocks/utils.py:71-189``, pinned by
``tools/make_vis_golden.py``) and ``cvpmvsnet`` (``blocks/cvp_mvsnet_components.py:192-245, :375-457``, pinned by
``tools/make_cvp_golden.py``), and ``mvs_variance_backward`` / ``mvs_regress_backward`` (autograd through ``mvsnet.py:123-141``,
pinned by ``tools/make_mvs_bwd_golden.py``).

Who may import it: ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` -- there only as the
checker or as the timed CPU baseline.  The product package ``robustmvd_b200``
never imports it and has no CPU fallback.

Parity pin: the reference ships no golden vectors or known-answer tests for this
path, so the oracle is pinned against the reference code
ITSELF, executed in the build container on CPU (``tools/validate_oracle.py``) and
frozen into ``tests/golden/*.npz`` by ``tools/make_golden.py``:
masks and integer sampling indices agree bit for bit, floating-point volumes
within 1e-5 of max|ref|.  Third-party arithmetic that the reference calls and
that is restated here: ``torch.nn.functional.grid_sample`` (bilinear, zeros
padding, ``align_corners=False``; torch 2.11 ATen CPU kernel, whose
unnormalisation is the fused ``fma(x + 1, size / 2, -0.5)``), ``torch.matmul``,
``F.softmax``, ``F.avg_pool3d``, ``torch.gather`` and
``kornia.utils.create_meshgrid`` (unpinned in the reference's requirements).
"""
from .planesweep import (  # noqa: F401
    compute_sampling_invdepths,
    epipolar_coeffs,
    sampling_points,
    grid_unnormalize,
    bilinear_setup,
    torch_corr,
    warp_only_corr,
    planesweep_correlation,
    planesweep_correlation_backward,
)
from .mvsnet import (  # noqa: F401
    homo_warp_coords,
    homo_warp,
    mvs_variance,
    mvs_regress,
    mvs_variance_backward,
    mvs_regress_backward,
    softargmax_partials,
    softargmax_combine,
    softargmax_confidence_part,
    soft_argmin,
    entropy,
)
from .fusion import view_fusion  # noqa: F401,E402
from .vismvsnet import homography_coords, homography_warp_volume, groupwise_cost_volume  # noqa: F401,E402
from . import cvpmvsnet  # noqa: F401,E402
