"""Solve time of mid-size bundles under the environment's PCG plan knobs (MPBA_COARSE_GROUPS, MPBA_PCG_MIN_CLUSTER)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mpsfm_b200 import engine, synthetic  # noqa: E402

for C, P, L in ((30, 15000, 5), (50, 30000, 5), (100, 50000, 6), (200, 100000, 6)):
    ba = engine.BundleAdjuster(synthetic.make_scene(C, P, L, seed=0))
    ba.solve()
    best = 1e9
    for _ in range(3):
        ba.reset_state()
        s = ba.solve()
        best = min(best, s.solve_ms)
    print(f"G={os.environ.get('MPBA_COARSE_GROUPS','-'):>2} cl>={os.environ.get('MPBA_PCG_MIN_CLUSTER','-'):>2}  C={C:4d} its {s.num_iterations:3d} pcg {s.total_pcg_iterations:5d}  solve {best:7.2f} ms")
