"""One small bundle solved with the host LM loop, a few iterations: the target of `ncu` launch lists.  GPU box only.
    python scripts/small_solve.py IMAGES POINTS OBS_PER_POINT [ITERATIONS]"""
import sys

sys.path.insert(0, ".")
from mpsfm_b200 import engine  # noqa: E402
from mpsfm_b200.synthetic import make_scene  # noqa: E402

C, P, L = (int(v) for v in sys.argv[1:4])
iters = int(sys.argv[4]) if len(sys.argv) > 4 else 3
ba = engine.BundleAdjuster(make_scene(C, P, L, seed=0))
s = ba.solve(engine.default_options(lm_loop=1, max_num_iterations=iters))
print(s.brief_report())
