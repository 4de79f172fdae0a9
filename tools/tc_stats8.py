"""Phase counters of the tcgen05 plane-sweep kernel in steady state (cfg2, 8 frames per launch): RMVD_PSWEEP_STATS=1."""
import os, sys, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from robustmvd_b200 import synthetic as syn
from robustmvd_b200.blocks.planesweep_corr import PlanesweepCorrelation
cu = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
frames = int(sys.argv[1]) if len(sys.argv) > 1 else 8
inp = syn.planesweep_inputs(48, 160, 256, 2, n=frames)
kw = dict(feat_key=cu(inp["feat_key"]), intrinsics_key=cu(inp["intrinsics_key"]),
          feat_sources=[cu(f) for f in inp["feat_sources"]], source_to_key_transforms=[cu(t) for t in inp["transforms"]],
          num_sampling_points=256, min_depth=0.4, max_depth=1000.0)
blk = PlanesweepCorrelation()
for _ in range(3):
    blk(**kw)
torch.cuda.synchronize()
os.environ["RMVD_PSWEEP_STATS"] = "1"
blk(**kw)
torch.cuda.synchronize()
