"""Host-buffer pipeline around ``PlanesweepCorrelation``: the call a host-side user makes when features arrive in (pinned)
host memory and the cost volumes are wanted back on the host.

The batch is cut into chunks of frames; the H2D copy of chunk i+1, the kernels of chunk i and the D2H copy of chunk i-1 run
on three CUDA streams, so the two PCIe directions (full duplex) and the GPU overlap.  Results are the block's own outputs,
bit for bit: every chunk is one ordinary ``PlanesweepCorrelation.forward`` call (reference signature,
rmvd/models/blocks/planesweep_corr.py:396-427) on a slice of the batch.
"""
import torch

from .blocks.planesweep_corr import PlanesweepCorrelation


class HostPlanesweepPipeline:
    def __init__(self, device, block=None, chunk_frames=2):
        self.device = torch.device(device)
        self.block = block if block is not None else PlanesweepCorrelation()
        # int: equal chunks; sequence: chunk sizes in order, the last one repeating (a small first chunk starts the D2H
        # stream earlier, which is the longer of the two copy directions)
        self.chunk = [int(c) for c in chunk_frames] if isinstance(chunk_frames, (list, tuple)) else [int(chunk_frames)]
        self.s_in = torch.cuda.Stream(self.device)
        self.s_cmp = torch.cuda.Stream(self.device)
        self.s_out = torch.cuda.Stream(self.device)
        self._dev_in = None   # persistent device copies of the inputs (allocated on first use, reused by later calls)

    def _inputs(self, host_key, host_sources):
        shapes = (tuple(host_key.shape), tuple(tuple(f.shape) for f in host_sources))
        if self._dev_in is None or self._dev_in[0] != shapes:
            key = torch.empty(host_key.shape, dtype=torch.float32, device=self.device)
            srcs = [torch.empty(f.shape, dtype=torch.float32, device=self.device) for f in host_sources]
            self._dev_in = (shapes, key, srcs)
        return self._dev_in[1], self._dev_in[2]

    @torch.no_grad()
    def run(self, host_key, intrinsics_key, host_sources, source_to_key_transforms, out_corr, out_mask, **kw):
        """``host_key [N,C,h,w]`` and ``host_sources`` (list of ``[N,C,hs,ws]``) are pinned host tensors,
        ``intrinsics_key [N,3,3]`` / ``source_to_key_transforms`` (list of ``[N,4,4]``) device tensors, ``out_corr`` /
        ``out_mask`` pinned host tensors ``[V,N,S,h,w]`` that receive the block's outputs.  ``kw``: the block's sampling
        arguments (``num_sampling_points, min_depth, max_depth`` ...).  Returns after everything has landed on the host."""
        N = host_key.shape[0]
        V = len(host_sources)
        d_key, d_src = self._inputs(host_key, host_sources)
        cur = torch.cuda.current_stream(self.device)
        for s in (self.s_in, self.s_cmp, self.s_out):
            s.wait_stream(cur)
        keep = []
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
            with torch.cuda.stream(self.s_cmp):
                self.s_cmp.wait_event(ev_in)
                corrs, masks, _ = self.block(feat_key=d_key[n0:n1], intrinsics_key=intrinsics_key[n0:n1],
                                             feat_sources=[f[n0:n1] for f in d_src],
                                             source_to_key_transforms=[t[n0:n1] for t in source_to_key_transforms], **kw)
                ev_cmp.record(self.s_cmp)
            with torch.cuda.stream(self.s_out):
                self.s_out.wait_event(ev_cmp)
                for v in range(V):
                    corrs[v].record_stream(self.s_out)
                    masks[v].record_stream(self.s_out)
                    out_corr[v, n0:n1].copy_(corrs[v], non_blocking=True)
                    out_mask[v, n0:n1].copy_(masks[v], non_blocking=True)
            keep.append((corrs, masks))
        self.s_out.synchronize()
        cur.wait_stream(self.s_cmp)
        return out_corr, out_mask
