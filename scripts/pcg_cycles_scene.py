"""PCG phase cycle counters (options.verbose = 2; the cluster kernel's detail needs a -DMPBA_PCG_TIMING build) for a
synthetic scene.
    [MPBA_LIB=build/libmpba_timing.so] MPBA_PLAN_DEBUG=1 python scripts/pcg_cycles_scene.py IMAGES POINTS OBS_PER_POINT [PCG_KERNEL]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mpsfm_b200 import engine, synthetic  # noqa: E402

C, P, L = (int(v) for v in sys.argv[1:4])
kern = int(sys.argv[4]) if len(sys.argv) > 4 else 2
ba = engine.BundleAdjuster(synthetic.make_scene(C, P, L, seed=0))
s = ba.solve(engine.default_options(max_num_iterations=3, verbose=2, profile=1, pcg_kernel=kern))
print(s.num_iterations, s.total_pcg_iterations, s.final_cost)
print({k: (round(1e3 * v[0] / v[1], 1), v[1]) for k, v in ba.kernel_times().items()})
