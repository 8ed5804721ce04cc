"""Does the upload of the next problem overlap the solve of the current one?  (cfg3, two pinned adjusters, copy stream.)"""
import sys, time
sys.path.insert(0, ".")
import torch
from mpsfm_b200 import engine
from mpsfm_b200.synthetic import make_config

prob = make_config("cfg3", seed=0)
lay = engine.build_layout(prob)
a = engine.BundleAdjuster(prob, layout=lay, pinned=True)
b = engine.BundleAdjuster(prob, layout=lay, pinned=True)
opt = engine.default_options()
for x in (a, b):
    x.upload(); x.solve(opt); x.state(copy=False)
cs = torch.cuda.Stream()
torch.cuda.synchronize()
t0 = time.perf_counter()
with torch.cuda.stream(cs):
    b.upload()
cs.synchronize()
print(f"upload alone {1e3*(time.perf_counter()-t0):.2f} ms")
a.upload(); torch.cuda.synchronize()
t0 = time.perf_counter(); a.solve(opt); print(f"solve alone {1e3*(time.perf_counter()-t0):.2f} ms")
a.upload(); torch.cuda.synchronize()
t0 = time.perf_counter()
with torch.cuda.stream(cs):
    b.upload()
a.solve(opt)
t2 = time.perf_counter()
cs.synchronize()
print(f"upload of b during solve of a: solve returned at {1e3*(t2-t0):.2f} ms, copy stream done at {1e3*(time.perf_counter()-t0):.2f} ms")
