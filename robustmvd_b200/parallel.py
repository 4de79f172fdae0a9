"""Depth-slab partition of the cost-volume path across the GPUs of one box, one process per GPU with
``torch.distributed`` (NCCL over NVLink / NVSwitch on the B200 box; the host logic is backend-agnostic and is tested
with ``gloo`` at world size 2 on CPU).

Every (sample, view, plane, pixel) output of the path is independent, so rank g computes planes
``slab_range(S, G, g)`` with the single-GPU kernels and

* ``planesweep_correlation_slabs``: all-gathers ``corr`` / ``mask`` along the plane axis (the consumer,
  ``LearnedFusion``, reads the planes as channels, learned_fusion.py:13-17);
* ``mvs_variance_slabs``: leaves the variance volume sharded along D (or gathers it on request);
* ``mvs_regress_slabs``: per-slab online-softmax partials (m, Z, A, I) -> all-gather (16 B / pixel / rank) -> merge ->
  each rank's share of the 4-bin confidence around the global index -> all-reduce(sum).

The reference has no equivalent: its only multi-GPU mode is ``nn.DataParallel`` over the batch (helpers.py:161-170).
The local compute steps default to the CUDA library; there is no CPU implementation in this package (tests inject the
oracle as the local compute to exercise the collective plumbing on CPU).
"""
import torch
import torch.distributed as dist


def slab_range(num_planes, world_size, rank):
    """Contiguous balanced partition: the first ``num_planes % world_size`` ranks get one extra plane."""
    base, extra = divmod(int(num_planes), int(world_size))
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def _world(group):
    return dist.get_world_size(group), dist.get_rank(group)


def all_gather_planes(local, num_planes, dim, group=None):
    """Concatenate per-rank slabs (``local.shape[dim]`` = this rank's slab length, slabs as in ``slab_range``) along
    ``dim``.  Slabs may differ by one plane, so they are padded to the longest for the fixed-size collective."""
    G, rank = _world(group)
    if G == 1:
        return local
    longest = slab_range(num_planes, G, 0)[1]
    x = local.movedim(dim, 0)
    if x.shape[0] < longest:
        pad = torch.zeros((longest - x.shape[0],) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
        x = torch.cat([x, pad], 0)
    x = x.contiguous()
    buf = torch.empty((G * longest,) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
    dist.all_gather_into_tensor(buf, x, group=group)   # concatenated form (accepted by NCCL and gloo alike)
    buf = buf.view((G, longest) + tuple(x.shape[1:]))
    parts = []
    for g in range(G):
        lo, hi = slab_range(num_planes, G, g)
        parts.append(buf[g, : hi - lo])
    return torch.cat(parts, 0).movedim(0, dim).contiguous()


def planesweep_correlation_slabs(block, group=None, gather=True, **kwargs):
    """``PlanesweepCorrelation.forward`` with the S planes partitioned over the ranks of ``group``.
    Returns ``(corrs, masks, sampling_invdepths)`` with full ``[N,S,h,w]`` volumes on every rank (``gather=True``) or
    this rank's slab (``gather=False``)."""
    G, rank = _world(group)
    inv = block.get_sampling_invdepths(kwargs.get("num_sampling_points"), kwargs.get("min_depth"),
                                       kwargs.get("max_depth"), kwargs.get("sampling_invdepths"),
                                       kwargs.get("sampling_type", "linear_invdepth"), device=kwargs["feat_key"].device)
    S = inv.shape[1]
    lo, hi = slab_range(S, G, rank)
    corrs, masks, inv = block(plane_range=(lo, hi), **kwargs)
    if not gather or G == 1:
        return corrs, masks, inv
    if corrs and block.warp_only is False:
        # one collective for all views and both volumes: [2, V, N, Sl, h, w]
        both = torch.stack([torch.stack(corrs, 0), torch.stack(masks, 0)], 0)
        full = all_gather_planes(both, S, 3, group)
        return list(full[0].unbind(0)), list(full[1].unbind(0)), inv
    corrs = [all_gather_planes(c, S, 1, group) for c in corrs]
    masks = [all_gather_planes(m, S, 1, group) for m in masks]
    return corrs, masks, inv


def mvs_variance_slabs(feats, proj_mats, depth_values, group=None, gather=False, variance_fn=None):
    """MVSNet variance volume with D partitioned: returns this rank's ``[B,C,Dl,h,w]`` slab and its ``(lo, hi)``,
    or the gathered full volume when ``gather=True``."""
    if variance_fn is None:
        from .blocks.utils import mvs_variance_volume as variance_fn
    G, rank = _world(group)
    D = depth_values.shape[1]
    lo, hi = slab_range(D, G, rank)
    slab = variance_fn(feats, proj_mats, depth_values, plane_range=(lo, hi))
    if gather:
        return all_gather_planes(slab, D, 2, group), (0, D)
    return slab, (lo, hi)


class _CudaSoftargmaxOps:
    @staticmethod
    def partial(cost_slab, depth_slab, d_begin):
        from .blocks.utils import softargmax_partial
        return softargmax_partial(cost_slab, depth_slab, d_begin)

    @staticmethod
    def combine(partials):
        from .blocks.utils import softargmax_combine
        return softargmax_combine(partials)

    @staticmethod
    def confidence(cost_slab, merged, D, d_begin):
        from .blocks.utils import softargmax_confidence
        return softargmax_confidence(cost_slab, merged, D, d_begin)


def mvs_regress_slabs(cost_slab, depth_slab, d_begin, num_planes, group=None, ops=None):
    """softmax over D + expected depth + 4-bin confidence (mvsnet.py:140-160) when each rank holds planes
    ``[d_begin, d_begin + Dl)`` of the regularised cost.  Returns the same dict as ``mvs_regress`` on every rank."""
    ops = ops or _CudaSoftargmaxOps
    G, rank = _world(group)
    part = ops.partial(cost_slab, depth_slab, d_begin)                         # [4,B,h,w]
    if G > 1:
        parts = torch.empty((G * 4,) + tuple(part.shape[1:]), dtype=part.dtype, device=part.device)
        dist.all_gather_into_tensor(parts, part.contiguous(), group=group)
        parts = parts.view((G,) + tuple(part.shape))
    else:
        parts = part[None]
    merged = ops.combine(parts)                                                # (m, Z, depth, expected index)
    conf = ops.confidence(cost_slab, merged, num_planes, d_begin)              # this slab's bins only
    if G > 1:
        dist.all_reduce(conf, op=dist.ReduceOp.SUM, group=group)
    eidx = merged[3]
    didx = eidx.to(torch.int64).clamp_(0, num_planes - 1)
    return dict(depth=merged[2], confidence=conf, depth_index=didx, expected_index=eidx)
