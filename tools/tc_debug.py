import os, sys, numpy as np, torch
sys.path.insert(0, '/root/repo')
from robustmvd_b200 import synthetic as syn
from robustmvd_b200.blocks.planesweep_corr import PlanesweepCorrelation
from tools import cases
cu = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
def run(inp, S, plane_range=None, **kw):
    return PlanesweepCorrelation()(feat_key=cu(inp["feat_key"]), intrinsics_key=cu(inp["intrinsics_key"]),
        feat_sources=[cu(f) for f in inp["feat_sources"]], source_to_key_transforms=[cu(t) for t in inp["transforms"]],
        num_sampling_points=S, min_depth=0.4, max_depth=1000.0, plane_range=plane_range)
os.environ["RMVD_PSWEEP_STATS"] = "1"
inp = syn.planesweep_inputs(48, 160, 256, 2)
print("cfg2 1 frame"); c, m, _ = run(inp, 256); torch.cuda.synchronize()
full = [x.clone() for x in c]
c2, _, _ = run(inp, 256, plane_range=(0, 37)); torch.cuda.synchronize()
d = (c2[0] - full[0][:, :37]).abs()
print("slab vs full max diff", d.max().item(), "nonzero", (d > 0).sum().item(), "max|corr|", full[0].abs().max().item())
os.environ["RMVD_PSWEEP_KERNEL"] = "band"
cb, mb, _ = run(inp, 256); torch.cuda.synchronize()
print("tc vs band max diff", (cb[0] - full[0]).abs().max().item(), "masks equal", torch.equal(mb[0], m[0]))
del os.environ["RMVD_PSWEEP_KERNEL"]
inp4 = syn.planesweep_inputs(96, 144, 256, 2, seed=77)
print("cfg4"); run(inp4, 256); torch.cuda.synchronize()
