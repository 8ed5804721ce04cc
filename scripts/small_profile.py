"""Per-kernel CUDA-event times of small bundles (host LM loop, options.profile = 1).  GPU box only."""
import sys

sys.path.insert(0, ".")
from mpsfm_b200 import engine  # noqa: E402
from mpsfm_b200.synthetic import make_scene  # noqa: E402

for C, P, L in ((6, 3000, 4), (20, 10000, 5), (50, 30000, 5), (100, 50000, 6)):
    prob = make_scene(C, P, L, seed=0)
    ba = engine.BundleAdjuster(prob)
    ba.solve(engine.default_options(lm_loop=1))
    ba.reset_state()
    s = ba.solve(engine.default_options(lm_loop=1, profile=1))
    kt = ba.kernel_times()
    tot = sum(v[0] for v in kt.values())
    print(f"{C} images, {s.num_iterations} LM iterations, {s.total_pcg_iterations} PCG iterations, kernels {1e3 * tot / s.num_iterations:.1f} us / iteration")
    for k, (ms, n) in sorted(kt.items(), key=lambda kv: -kv[1][0]):
        print(f"   {k:28s} {1e3 * ms / n:8.2f} us x {n / s.num_iterations:.1f} / iteration")
