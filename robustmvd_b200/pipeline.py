"""Host-buffer pipeline around ``PlanesweepCorrelation``: the call a host-side user makes when features arrive in (pinned)
host memory and the cost volumes are wanted back on the host.

The batch is cut into chunks of frames; the H2D copy of chunk i+1, the kernels of chunk i and the D2H copy of chunk i-1 run
on three CUDA streams, so the two PCIe directions (full duplex) and the GPU overlap.  Results are the block's own outputs,
bit for bit: every chunk is one ordinary ``PlanesweepCorrelation.forward`` call (reference signature,
rmvd/models/blocks/planesweep_corr.py:396-427) on a slice of the batch.

A stream of batches (``run(..., wait=False)`` per batch, ``finish()`` at the end) also overlaps consecutive batches: the
device-side input buffers are double-buffered, so the H2D copy of batch i+1 runs under the kernels and the D2H copy of batch
i and the steady-state cost of a batch is the longer of its two PCIe directions.

``compact_mask=True`` (default) sends the sampling mask over PCIe as one byte per element instead of four: the mask holds only
0.0 and 1.0 (planesweep_corr.py:99-102, :189-191) and is half of the device-to-host bytes.  ``rmvd_mask_pack_u8`` packs it on the
device, the bytes land in a pinned staging ring, and ``rmvd_host_expand_mask_async`` -- a host function queued on a fourth stream,
running on the library's worker threads -- writes the float32 values into ``out_mask``; the host buffer ends up bit-identical to
the plain float32 copy (``compact_mask=False``).  ``mask_transport="bits"`` sends one BIT per element (``rmvd_mask_pack_bits`` /
``rmvd_host_expand_mask_bits_async``; per-view element counts that are not multiples of 32 fall back to bytes): an eighth of the
mask bytes again, but measured slower end to end on the B200 box (9.0-9.9 ms against 8.6-8.8 ms per 16-frame step), so bytes are
the default.
"""
import ctypes
import os

import torch

from . import _lib
from .blocks.planesweep_corr import PlanesweepCorrelation


def _default_mask_threads():
    try:
        cores = len(os.sched_getaffinity(0))
    except AttributeError:
        cores = os.cpu_count() or 1
    ranks = int(os.environ.get("LOCAL_WORLD_SIZE", os.environ.get("WORLD_SIZE", "1")) or 1)
    # measured on the B200 box (16 host cores, cfg2, 16 frames per step): 8.60 / 8.81 / 9.16 / 9.71 ms per step with 2 / 4 / 8 / 12
    # threads (more threads take memory bandwidth and cores from the DMA engines and the issuing Python thread), 2 is at the
    # edge of keeping up with the copies: 3
    return max(1, min(3, cores // max(1, ranks) - 1))


class HostPlanesweepPipeline:
    def __init__(self, device, block=None, chunk_frames=2, compact_mask=True, mask_threads=None, mask_transport="u8"):
        self.device = torch.device(device)
        self.compact_mask = bool(compact_mask)
        assert mask_transport in ("bits", "u8")
        self.mask_transport = mask_transport
        self.mask_threads = int(mask_threads) if mask_threads else _default_mask_threads()
        self.s_exp = torch.cuda.Stream(self.device)      # carries the host-side mask expansion (host functions only)
        self._stage = [None] * 4                          # pinned byte staging ring: [tensor, event after its expansion]
        self._stages = 0
        self.block = block if block is not None else PlanesweepCorrelation()
        # int: equal chunks; sequence: chunk sizes in order, the last one repeating (a small first chunk starts the D2H
        # stream earlier, which is the longer of the two copy directions)
        self.chunk = [int(c) for c in chunk_frames] if isinstance(chunk_frames, (list, tuple)) else [int(chunk_frames)]
        self.s_in = torch.cuda.Stream(self.device)
        self.s_cmp = torch.cuda.Stream(self.device)
        self.s_out = torch.cuda.Stream(self.device)
        # persistent device copies of the inputs, two sets used alternately by consecutive calls (allocated on first use)
        self._dev_in = [None, None]
        self._cmp_done = [None, None]   # event after the last kernel that read each set
        self._calls = 0
        self._done = None               # event after the last D2H copy of the latest call
        # persistent device outputs: a ring of chunk-sized (corr, mask) pairs, each reusable once its D2H copy has completed
        self._out_ring = [None] * 4
        self._chunks = 0

    def _inputs(self, host_key, host_sources, slot):
        shapes = (tuple(host_key.shape), tuple(tuple(f.shape) for f in host_sources))
        if self._dev_in[slot] is None or self._dev_in[slot][0] != shapes:
            if self._cmp_done[slot] is not None:
                self._cmp_done[slot].synchronize()      # the old buffers of this set may still be in use
            key = torch.empty(host_key.shape, dtype=torch.float32, device=self.device)
            srcs = [torch.empty(f.shape, dtype=torch.float32, device=self.device) for f in host_sources]
            self._dev_in[slot] = (shapes, key, srcs)
        return self._dev_in[slot][1], self._dev_in[slot][2]

    def _outputs(self, shape):
        slot = self._chunks % len(self._out_ring)
        self._chunks += 1
        e = self._out_ring[slot]
        if e is None or e[0] != shape:
            if e is not None:
                e[3].synchronize()
            e = [shape, torch.empty(shape, dtype=torch.float32, device=self.device),
                 torch.empty(shape, dtype=torch.float32, device=self.device), None,
                 torch.empty(shape, dtype=torch.uint8, device=self.device) if self.compact_mask else None]
            self._out_ring[slot] = e
        elif e[3] is not None:
            self.s_cmp.wait_event(e[3])                   # the previous contents have been copied out
        return e

    def _staging(self, shape):
        """Next pinned byte buffer of the staging ring; ``s_out`` waits until the expansion that last read it has run."""
        slot = self._stages % len(self._stage)
        self._stages += 1
        e = self._stage[slot]
        if e is None or tuple(e[0].shape) != tuple(shape):
            if e is not None and e[1] is not None:
                e[1].synchronize()
            e = [torch.empty(shape, dtype=torch.uint8).pin_memory(), None]
            self._stage[slot] = e
        elif e[1] is not None:
            self.s_out.wait_event(e[1])
        return e

    def finish(self):
        """Block the host until every batch handed to ``run`` has landed in its host buffers; the current stream is ordered
        after them as well."""
        if self._done is not None:
            torch.cuda.current_stream(self.device).wait_event(self._done)
            self._done.synchronize()

    @torch.no_grad()
    def run(self, host_key, intrinsics_key, host_sources, source_to_key_transforms, out_corr, out_mask, wait=True, **kw):
        """``host_key [N,C,h,w]`` and ``host_sources`` (list of ``[N,C,hs,ws]``) are pinned host tensors,
        ``intrinsics_key [N,3,3]`` / ``source_to_key_transforms`` (list of ``[N,4,4]``) device tensors, ``out_corr`` /
        ``out_mask`` pinned host tensors ``[V,N,S,h,w]`` that receive the block's outputs.  ``kw``: the block's sampling
        arguments (``num_sampling_points, min_depth, max_depth`` ...).  Returns after everything has landed on the host;
        with ``wait=False`` it returns once the work is queued (the host buffers must stay untouched until ``finish()``)."""
        N = host_key.shape[0]
        V = len(host_sources)
        lib = _lib.load()
        if self.compact_mask and not (out_mask.is_contiguous() and out_mask.dtype == torch.float32 and not out_mask.is_cuda):
            raise _lib.RmvdError("compact_mask needs a contiguous float32 host tensor for out_mask")
        slot = self._calls & 1
        self._calls += 1
        d_key, d_src = self._inputs(host_key, host_sources, slot)
        cur = torch.cuda.current_stream(self.device)
        for s in (self.s_in, self.s_cmp, self.s_out, self.s_exp):
            s.wait_stream(cur)
        if self._cmp_done[slot] is not None:              # kernels of the call before last read this set of input buffers
            self.s_in.wait_event(self._cmp_done[slot])
        bounds, n0 = [], 0
        while n0 < N:
            c = self.chunk[min(len(bounds), len(self.chunk) - 1)]
            bounds.append((n0, min(N, n0 + c)))
            n0 += c
        for n0, n1 in bounds:
            ev_in, ev_cmp = torch.cuda.Event(), torch.cuda.Event()
            with torch.cuda.stream(self.s_in):
                d_key[n0:n1].copy_(host_key[n0:n1], non_blocking=True)
                for v in range(V):
                    d_src[v][n0:n1].copy_(host_sources[v][n0:n1], non_blocking=True)
                ev_in.record(self.s_in)
            S = kw.get("num_sampling_points")
            pr = kw.get("plane_range")
            if S is None:
                S = kw["sampling_invdepths"].shape[1]
            planes = S if pr is None else pr[1] - pr[0]
            ring = self._outputs((V, n1 - n0, planes, host_key.shape[2], host_key.shape[3]))
            with torch.cuda.stream(self.s_cmp):
                self.s_cmp.wait_event(ev_in)
                corrs, masks, _ = self.block(feat_key=d_key[n0:n1], intrinsics_key=intrinsics_key[n0:n1],
                                             feat_sources=[f[n0:n1] for f in d_src],
                                             source_to_key_transforms=[t[n0:n1] for t in source_to_key_transforms],
                                             out=(ring[1], ring[2]), **kw)
                per_view = ring[2][0].numel()
                bits = self.compact_mask and self.mask_transport == "bits" and per_view % 32 == 0
                if bits:        # view v's words start at byte v * per_view / 8 of the chunk's byte buffer (ring[4] is sized for bytes)
                    _lib.check(lib.rmvd_mask_pack_bits(_lib.ptr(ring[2]), _lib.ptr(ring[4]), ring[2].numel(), _lib.stream_ptr()),
                               "rmvd_mask_pack_bits")
                elif self.compact_mask:
                    _lib.check(lib.rmvd_mask_pack_u8(_lib.ptr(ring[2]), _lib.ptr(ring[4]), ring[2].numel(), _lib.stream_ptr()),
                               "rmvd_mask_pack_u8")
                ev_cmp.record(self.s_cmp)
            with torch.cuda.stream(self.s_out):
                self.s_out.wait_event(ev_cmp)
                stage = self._staging(ring[4].shape) if self.compact_mask else None
                for v in range(V):                        # one contiguous block per view
                    out_corr[v, n0:n1].copy_(ring[1][v], non_blocking=True)
                    if not self.compact_mask:
                        out_mask[v, n0:n1].copy_(ring[2][v], non_blocking=True)
                if bits:
                    nb = ring[2].numel() // 8
                    stage[0].view(-1)[:nb].copy_(ring[4].view(-1)[:nb], non_blocking=True)
                elif self.compact_mask:
                    stage[0].copy_(ring[4], non_blocking=True)
                ring[3] = torch.cuda.Event()
                ring[3].record(self.s_out)
            if self.compact_mask:
                with torch.cuda.stream(self.s_exp):
                    self.s_exp.wait_event(ring[3])
                    for v in range(V):
                        dst = out_mask[v, n0:n1]
                        if bits:
                            _lib.check(lib.rmvd_host_expand_mask_bits_async(
                                ctypes.c_void_p(stage[0].data_ptr() + v * (per_view // 8)), ctypes.c_void_p(dst.data_ptr()), dst.numel(),
                                self.mask_threads, _lib.stream_ptr()), "rmvd_host_expand_mask_bits_async")
                        else:
                            _lib.check(lib.rmvd_host_expand_mask_async(ctypes.c_void_p(stage[0][v].data_ptr()), ctypes.c_void_p(dst.data_ptr()),
                                                                       dst.numel(), self.mask_threads, _lib.stream_ptr()),
                                       "rmvd_host_expand_mask_async")
                    stage[1] = torch.cuda.Event()
                    stage[1].record(self.s_exp)
            self._cmp_done[slot] = ev_cmp
        self._done = torch.cuda.Event()
        if self.compact_mask:
            self.s_exp.wait_stream(self.s_out)            # (already implied by the last chunk's event; keeps the intent explicit)
            self._done.record(self.s_exp)
        else:
            self._done.record(self.s_out)
        if wait:
            self.finish()
        return out_corr, out_mask


class GraphedPlanesweep:
    """One ``PlanesweepCorrelation.forward`` of a fixed shape, captured in a CUDA graph and replayed: the launch sequence (two
    split-copy kernels + the persistent plane-sweep kernel) costs no Python, no ctypes marshalling and no allocation per call.
    That is the latency mode for the batch size the reference evaluates at (one frame per call,
    rmvd/eval/multi_view_depth_evaluation.py:234-240): 0.061 ms per KITTI-shape frame against 0.093 ms through the eager block.

    ``example`` = the keyword arguments of one representative call (device tensors).  ``__call__(**kwargs)`` copies the tensors
    it is given into the captured input buffers (device-to-device, stream-ordered) and replays; it returns the block's outputs,
    which live in the graph's own memory pool and are overwritten by the next call."""

    TENSOR_KEYS = ("feat_key", "intrinsics_key", "feat_sources", "source_to_key_transforms", "intrinsics_sources", "sampling_invdepths")

    def __init__(self, example, block=None):
        self.block = block if block is not None else PlanesweepCorrelation()
        clone = lambda v: [t.clone() for t in v] if isinstance(v, (list, tuple)) else v.clone()  # noqa: E731
        self.static = {k: (clone(v) if k in self.TENSOR_KEYS and v is not None else v) for k, v in example.items()}
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side), torch.no_grad():          # warm-up outside the capture (library load, workspace, schedule cache)
            self.block(**self.static)
        torch.cuda.current_stream().wait_stream(side)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph), torch.no_grad():
            self.out = self.block(**self.static)

    @torch.no_grad()
    def __call__(self, **kwargs):
        for k, v in kwargs.items():
            if k not in self.TENSOR_KEYS or v is None:
                if k not in self.TENSOR_KEYS and self.static.get(k) != v:
                    raise ValueError(f"GraphedPlanesweep was captured with {k}={self.static.get(k)!r}")
                continue
            dst = self.static[k]
            if isinstance(v, (list, tuple)):
                for d, s in zip(dst, v):
                    if d.data_ptr() != s.data_ptr():
                        d.copy_(s, non_blocking=True)
            elif dst.data_ptr() != v.data_ptr():
                dst.copy_(v, non_blocking=True)
        self.graph.replay()
        return self.out
