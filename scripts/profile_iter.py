"""A few LM iterations of one workload, for ncu: python scripts/profile_iter.py [--workload cfg3] [--iters 2]"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from mpsfm_b200 import engine  # noqa: E402
from mpsfm_b200.synthetic import make_config  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="cfg3")
ap.add_argument("--iters", type=int, default=2)
a = ap.parse_args()
ba = engine.BundleAdjuster(make_config(a.workload, seed=0), device=0)
s = ba.solve(engine.default_options(max_num_iterations=a.iters, profile=1))
print(s.brief_report(), "launches", s.kernel_launches)
print({k: round(1e3 * ms / n, 1) for k, (ms, n) in ba.kernel_times().items()})
