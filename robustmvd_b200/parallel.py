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


class SlabGather:
    """Depth slabs with the gather fused into the plane-sweep kernel (partition 1).

    Every rank holds symmetric ``[2, V, N, S, h, w]`` output buffers (``torch.distributed._symmetric_memory``: peer-mapped over
    NVLink / NVSwitch).  A call computes this rank's planes ``slab_range(S, G, rank)`` and its sampling loop stores each corr / mask
    row into the buffers of ALL ranks; the kernel ends with a flag handshake between the ranks (``rmvd_psweep_corr_fwd_peers``), so
    one launch sequence per rank replaces "slab kernel + stack / movedim / contiguous + NCCL all-gather + cat".  Two buffer sets
    alternate between consecutive calls: a rank that runs ahead writes set ``(k + 1) % 2`` while a slower rank still reads set
    ``k % 2``, and cannot start call ``k + 2`` before the slower rank has finished call ``k + 1`` (the handshake).

    ``transport="multicast"``: the sampling loop issues one ``multimem.st`` per value to the buffers' NVSwitch multicast address (NVLS)
    instead of one store per rank; ``transport="copy"``: see ``_call_copy``; ``transport="push"``: the sweep stores locally and a pusher
    kernel on ``push_ctas`` SMs of a side stream forwards the finished chunks (``rmvd_psweep_corr_fwd_push``).  ``partition="tiles"``: the ranks split the (tile, sample, view) work items instead of the planes -- same result on every rank, and
    the per-tile costs that do not shrink with the plane count (key tile, band planning, band rows) are divided too; at one frame
    this is the partition that is faster than one GPU.

    Returns ``(corrs, masks, sampling_invdepths)`` with full ``[N,S,h,w]`` volumes per view, views of the current buffer set: valid
    until the call after next.  Falls back to ``planesweep_correlation_slabs`` (NCCL / gloo all-gather) when symmetric memory is
    not available (``mode == "all_gather"``), e.g. in the CPU tests of the host logic.
    """

    def __init__(self, block, V, N, S, h, w, device, group=None, partition="planes", transport="stores", view_groups=None,
                 push_ctas=8):
        assert partition in ("planes", "tiles") and transport in ("stores", "multicast", "copy", "push")
        assert not (transport == "push" and partition != "planes")
        self.push_ctas = push_ctas
        self.block, self.group = block, group
        self.G, self.rank = _world(group)
        self.S = S
        self.partition = partition
        self.transport = transport
        self.view_groups = view_groups
        self.epoch = 0
        self.mode = "all_gather"
        self.bufs = None
        if torch.device(device).type == "cuda" and self.G > 1 and self.G <= 8:
            try:
                import torch.distributed._symmetric_memory as symm
                grp = group if group is not None else dist.group.WORLD
                self.bufs, self.hdls = [], []
                for _ in range(2):
                    t = symm.empty((2, V, N, S, h, w), dtype=torch.float32, device=device)
                    self.bufs.append(t)
                    self.hdls.append(symm.rendezvous(t, grp))
                self.flags = symm.empty((64,), dtype=torch.int32, device=device)
                self.flags.zero_()
                self.fh = symm.rendezvous(self.flags, grp)
                torch.cuda.synchronize(device)
                self.fh.barrier(channel=0)           # every rank's flags are zero before anyone's first epoch arrives
                torch.cuda.synchronize(device)
                self.mode = "peer_stores" if transport in ("stores", "multicast", "push") else "copy_engine"
                self.mc = [0, 0]
                if transport == "multicast":     # NVLS: one multimem.st per value, replicated by the switch into every rank's buffer
                    self.mc = [int(hd.multicast_ptr) for hd in self.hdls]
                    if not all(self.mc):
                        raise RuntimeError("no multicast address: NVLS is not available for this group")
                    self.mode = "peer_multicast"
                if transport == "push":          # a pusher kernel on a side stream forwards finished chunks to the peers
                    from . import _lib
                    lo, hi = slab_range(S, self.G, self.rank)
                    nb = _lib.load().rmvd_psweep_push_queue_bytes(N, h, w, V, max(hi - lo, 1))
                    self.queue = torch.zeros((nb + 15) // 16 * 4, dtype=torch.int32, device=device)
                    self.side = torch.cuda.Stream(device=device)
                    torch.cuda.synchronize(device)
                    self.mode = "peer_push"
                if transport == "copy":
                    self.side = torch.cuda.Stream(device=device)
                    shape = (2, V, N, S, h, w)
                    self.peer_bufs = [[hd.get_buffer(p, shape, torch.float32) for p in range(self.G)] for hd in self.hdls]
            except Exception as e:  # noqa: BLE001
                self.bufs = None
                self.why = repr(e)

    def _call_copy(self, **kwargs):
        """Depth slabs with the copy engines as the transport: the views are cut into groups, the slab of group k is computed by the
        single-GPU kernel straight into this rank's symmetric buffer (at its plane offset) and then pushed to every peer's buffer by
        one 2-D copy per peer on a side stream while the kernel of group k + 1 runs; a stream-ordered barrier between the ranks
        closes the call.  The persistent kernel holds every SM, the copy engines hold none: the two overlap, which SM-issued peer
        stores from the sampling loop do not (they stall the consumer warps, profiles/r02t)."""
        from . import _lib
        self.epoch += 1
        k = self.epoch % 2
        buf = self.bufs[k]
        lo, hi = slab_range(self.S, self.G, self.rank)
        srcs = list(kwargs.pop("feat_sources"))
        Ts = list(kwargs.pop("source_to_key_transforms"))
        Ks = kwargs.pop("intrinsics_sources", None)
        V = len(srcs)
        ng = self.view_groups or min(V, 4)
        bounds = [round(i * V / ng) for i in range(ng + 1)]
        main = torch.cuda.current_stream()
        lib = _lib.load()
        N, h, w = buf.shape[2], buf.shape[4], buf.shape[5]
        plane = h * w * 4
        pitch, row = self.S * plane, (hi - lo) * plane
        inv = None
        self.side.wait_stream(main)                  # the previous readers of this buffer set are ahead of us in stream order
        for gi in range(ng):
            v0, v1 = bounds[gi], bounds[gi + 1]
            if v1 == v0:
                continue
            if hi > lo:
                _, _, inv = self.block._forward_impl(
                    feat_sources=srcs[v0:v1], source_to_key_transforms=Ts[v0:v1], intrinsics_sources=None if Ks is None else Ks[v0:v1],
                    plane_range=(lo, hi), out=(buf[0, v0:v1], buf[1, v0:v1]), **kwargs)
            ev = torch.cuda.Event()
            ev.record(main)
            self.side.wait_event(ev)
            if hi > lo:
                for c in range(2):                   # corr, mask: rows = (view, sample) volumes of the group
                    off = ((c * V + v0) * N) * pitch + lo * plane
                    for p in range(self.G):
                        if p != self.rank:
                            st = lib.rmvd_copy_rows(self.peer_bufs[k][p].data_ptr() + off, buf.data_ptr() + off, pitch, row, (v1 - v0) * N,
                                                    self.side.cuda_stream)
                            _lib.check(st, "rmvd_copy_rows")
        with torch.cuda.stream(self.side):
            self.hdls[k].barrier(channel=1)          # every rank's copies into every buffer are complete
        main.wait_stream(self.side)
        if inv is None:
            inv = self.block.get_sampling_invdepths(kwargs.get("num_sampling_points"), kwargs.get("min_depth"), kwargs.get("max_depth"),
                                                    kwargs.get("sampling_invdepths"), kwargs.get("sampling_type", "linear_invdepth"),
                                                    device=buf.device)
        return list(buf[0].unbind(0)), list(buf[1].unbind(0)), inv

    def __call__(self, **kwargs):
        if self.mode == "copy_engine":
            return self._call_copy(**kwargs)
        if self.mode not in ("peer_stores", "peer_multicast", "peer_push"):
            return planesweep_correlation_slabs(self.block, self.group, **kwargs)
        self.epoch += 1
        k = self.epoch % 2
        buf, hdl = self.bufs[k], self.hdls[k]
        half = buf[0].numel() * 4                   # bytes of the corr half of a buffer; the mask half follows it
        peers = dict(rank=self.rank, epoch=self.epoch, corr=[int(p) for p in hdl.buffer_ptrs],
                     mask=[int(p) + half for p in hdl.buffer_ptrs], flags=[int(p) for p in self.fh.buffer_ptrs])
        if self.mode == "peer_multicast":
            peers.update(corr_mc=self.mc[k], mask_mc=self.mc[k] + half)
        lo, hi = slab_range(self.S, self.G, self.rank)
        if self.partition == "tiles":               # every rank sweeps all planes of every G-th tile (RMVD_PARTITION_TILES)
            lo, hi = 0, self.S
            peers["partition"] = 1
        if self.mode == "peer_push":
            main = torch.cuda.current_stream()
            peers.update(queue=self.queue, push_stream=self.side, push_ctas=self.push_ctas)
            _, _, inv = self.block._forward_impl(plane_range=(lo, hi), peers=peers, **kwargs)
            main.wait_stream(self.side)             # the pusher's handshake closes the call; it also guards the queue's reuse
            return list(buf[0].unbind(0)), list(buf[1].unbind(0)), inv
        _, _, inv = self.block._forward_impl(plane_range=(lo, hi), peers=peers, **kwargs)
        return list(buf[0].unbind(0)), list(buf[1].unbind(0)), inv


class SlabRegress:
    """``mvs_regress`` with D sharded over the ranks as ONE kernel per rank (``rmvd_softargmax_slab_fused``): partials into every
    rank's buffer by peer stores, in-kernel handshake, merge, confidence from peer loads -- no NCCL call.  ``self.cost`` is this
    rank's symmetric slab buffer ``[B, Dl_max, h, w]``: an upstream producer may write its slab there directly; ``__call__`` copies
    a given ``cost_slab`` into it first.  Falls back to ``mvs_regress_slabs`` when symmetric memory is unavailable."""

    def __init__(self, B, D, h, w, device, group=None):
        self.group = group
        self.G, self.rank = _world(group)
        self.B, self.D, self.h, self.w = B, D, h, w
        self.lo, self.hi = slab_range(D, self.G, self.rank)
        self.epoch = 0
        self.mode = "all_gather"
        if torch.device(device).type == "cuda" and 1 < self.G <= 8:
            try:
                import torch.distributed._symmetric_memory as symm
                grp = group if group is not None else dist.group.WORLD
                dmax = -(-D // self.G)
                self.cost = symm.empty((B, dmax, h, w), dtype=torch.float32, device=device)
                self.parts = symm.empty((self.G, 4, B, h, w), dtype=torch.float32, device=device)
                self.flags = symm.empty((64,), dtype=torch.int32, device=device)
                self.flags.zero_()
                self.hc, self.hp, self.hf = (symm.rendezvous(t, grp) for t in (self.cost, self.parts, self.flags))
                self.counters = torch.zeros(2, dtype=torch.int32, device=device)
                torch.cuda.synchronize(device)
                self.hf.barrier(channel=0)
                torch.cuda.synchronize(device)
                self.mode = "fused_kernel"
            except Exception as e:  # noqa: BLE001
                self.why = repr(e)

    def __call__(self, cost_slab, depth_values):
        """cost_slab ``[B, hi - lo, h, w]`` (or None: already in ``self.cost``), depth_values ``[B, D]`` (all planes)."""
        if self.mode != "fused_kernel":
            return mvs_regress_slabs(cost_slab, depth_values[:, self.lo:self.hi].contiguous(), self.lo, self.D, self.group)
        from . import _lib
        lib = _lib.load()
        dev = self.cost.device
        if cost_slab is not None:
            self.cost[:, : self.hi - self.lo].copy_(cost_slab)
        self.epoch += 1
        depth = torch.empty((self.B, self.h, self.w), dtype=torch.float32, device=dev)
        conf = torch.empty_like(depth)
        eidx = torch.empty_like(depth)
        didx = torch.empty((self.B, self.h, self.w), dtype=torch.int64, device=dev)
        dv = depth_values.detach().to(torch.float32).contiguous()
        st = lib.rmvd_softargmax_slab_fused(
            _lib.raw_ptr_array(self.hc.buffer_ptrs), _lib.ptr(dv), self.B, self.D, self.h, self.w, self.G, self.rank,
            _lib.raw_ptr_array(self.hp.buffer_ptrs), _lib.raw_ptr_array(self.hf.buffer_ptrs), self.epoch, _lib.ptr(self.counters),
            _lib.ptr(depth), _lib.ptr(conf), _lib.ptr(didx), _lib.ptr(eidx), _lib.stream_ptr())
        _lib.check(st, "rmvd_softargmax_slab_fused")
        return dict(depth=depth, confidence=conf, depth_index=didx, expected_index=eidx)


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
