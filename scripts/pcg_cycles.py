"""Prints the PCG kernel's phase cycle counters (mpba_options.verbose = 2; the cluster kernel's detail needs a library
built with -DMPBA_PCG_TIMING, e.g. MPBA_LIB=build/libmpba_timing.so) for a few LM iterations of a workload.

    python scripts/pcg_cycles.py cfg3
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mpsfm_b200 import engine, synthetic  # noqa: E402

w = sys.argv[1] if len(sys.argv) > 1 else "target10m"
ba = engine.BundleAdjuster(synthetic.make_config(w, seed=0))
s = ba.solve(engine.default_options(max_num_iterations=4, verbose=2, profile=1))
print(s.num_iterations, s.total_pcg_iterations, s.final_cost)
print({k: (round(1e3 * v[0] / v[1], 1), v[1]) for k, v in ba.kernel_times().items()})
