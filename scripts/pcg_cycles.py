"""Prints the PCG kernel's phase cycle counters (mpba_options.verbose = 2) for a few LM iterations of a workload.

    python scripts/pcg_cycles.py target10m
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mpsfm_b200 import engine, synthetic  # noqa: E402
from mpsfm_b200.problem import CameraModel  # noqa: E402

w = sys.argv[1] if len(sys.argv) > 1 else "target10m"
C, P, L = synthetic.CONFIGS[w]
prob = synthetic.make_scene(C, P, L, seed=0, model=CameraModel.SIMPLE_RADIAL)
ba = engine.BundleAdjuster(prob)
s = ba.solve(engine.default_options(max_num_iterations=4, verbose=2, profile=1))
print(s.num_iterations, s.total_pcg_iterations, s.final_cost)
print({k: (round(v[0], 3), v[1]) for k, v in ba.kernel_times().items()})
