"""Soak test of the tcgen05 plane-sweep kernel: many back-to-back 8-frame launches (hang / race detector) with a bit-exact
comparison of every launch against the first."""
import os, sys, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from robustmvd_b200 import synthetic as syn
from robustmvd_b200.blocks.planesweep_corr import PlanesweepCorrelation
cu = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 200
inp = syn.planesweep_inputs(48, 160, 256, 2, n=8)
kw = dict(feat_key=cu(inp["feat_key"]), intrinsics_key=cu(inp["intrinsics_key"]),
          feat_sources=[cu(f) for f in inp["feat_sources"]], source_to_key_transforms=[cu(t) for t in inp["transforms"]],
          num_sampling_points=256, min_depth=0.4, max_depth=1000.0)
blk = PlanesweepCorrelation()
c0, m0, _ = blk(**kw)
c0 = [x.clone() for x in c0]; m0 = [x.clone() for x in m0]
bad = 0
for r in range(reps):
    c, m, _ = blk(**kw)
    if r % 20 == 0:
        torch.cuda.synchronize()
        bad += sum(int(not torch.equal(a, b)) for a, b in zip(c, c0)) + sum(int(not torch.equal(a, b)) for a, b in zip(m, m0))
        print("rep", r, "mismatching tensors so far", bad, flush=True)
torch.cuda.synchronize()
print("soak done", reps, "launches, mismatches", bad)
