"""CPU oracle of the learned view fusion's elementwise part -- TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Restates rmvd/models/blocks/learned_fusion.py:24-54 from the per-view logits on: ``softmax(stack(logits), 0) + 1e-9`` (:32-33),
``view_weight * mask`` (:38-41), ``view_weights_sum`` (:42), ``fused_mask = (sum != 0)`` (:43), ``corr_sum`` (:45-47),
``corr_avg = corr_sum / (sum + 1e-9) * fused_mask`` (:48).  The convolutions that produce the logits (:12-16, :28) are
cuDNN / torch modules on both sides and not part of the oracle.  Pinned by ``tests/golden/fusion_*.npz``, frozen from the
unmodified reference module by ``tools/make_fusion_golden.py``.
"""
import numpy as np


def view_fusion(logits, corrs, masks):
    """logits: list of V arrays [N,1,h,w]; corrs, masks: lists of V arrays [N,S,h,w] (float32).
    Returns (view_weights [V,N,1,h,w], fused_corr [N,S,h,w], fused_mask [N,S,h,w]); V == 1 passes through (:50-52)."""
    V = len(corrs)
    if V == 1:
        return None, np.asarray(corrs[0], np.float32), np.asarray(masks[0], np.float32)
    x = np.stack([np.asarray(l, np.float32) for l in logits], 0)                     # :32
    e = np.exp(x - x.max(0, keepdims=True), dtype=np.float32)
    w = (e / e.sum(0, keepdims=True, dtype=np.float32)).astype(np.float32) + np.float32(1e-9)   # :33
    wm = [(w[v] * np.asarray(masks[v], np.float32)).astype(np.float32) for v in range(V)]      # :38-41
    wsum = np.zeros_like(wm[0])
    csum = np.zeros_like(wm[0])
    for v in range(V):                                                                # :42, :45-47 (sum over dim 0, in order)
        wsum = (wsum + wm[v]).astype(np.float32)
        csum = (csum + (np.asarray(corrs[v], np.float32) * wm[v]).astype(np.float32)).astype(np.float32)
    fmask = (wsum != 0).astype(np.float32)                                            # :43
    fcorr = ((csum / (wsum + np.float32(1e-9))).astype(np.float32) * fmask).astype(np.float32)   # :48
    return w, fcorr, fmask
