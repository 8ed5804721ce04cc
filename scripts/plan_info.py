"""Print the cluster plan of the PCG (MPBA_PLAN_DEBUG) and the first LM iterations of a workload (GPU box only)."""
import os
import sys

os.environ["MPBA_PLAN_DEBUG"] = "1"
sys.path.insert(0, ".")
from mpsfm_b200 import engine  # noqa: E402
from mpsfm_b200.synthetic import make_config  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "cfg3"
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 3
ba = engine.BundleAdjuster(make_config(name))
s = ba.solve(engine.default_options(max_num_iterations=iters, verbose=1))
print(s.brief_report())
