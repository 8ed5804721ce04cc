"""Print the PCG phase cycle counters (CTA 0) for a few LM iterations of a workload (GPU box only)."""
import sys

sys.path.insert(0, ".")
from mpsfm_b200 import engine
from mpsfm_b200.synthetic import make_config

name = sys.argv[1] if len(sys.argv) > 1 else "cfg3"
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 3
prob = make_config(name)
ba = engine.BundleAdjuster(prob)
s = ba.solve(engine.default_options(max_num_iterations=iters, verbose=2, profile=1))
print(s.brief_report())
print(ba.kernel_times())
