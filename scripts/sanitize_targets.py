"""Small runs that execute every kernel with hand-rolled synchronisation, sized for compute-sanitizer (10-100x slower):

    compute-sanitizer --tool racecheck python scripts/sanitize_targets.py cluster dense coarse4 lmgraph integrate
    compute-sanitizer --tool memcheck  python scripts/sanitize_targets.py all

  grid      k_pcg: cooperative grid, barrier = atomic arrive + spin on a monotonic counter (ba_kernels.cu)
  cluster   k_pcg_cluster: DSMEM deposits / halo reads between cluster barriers (pcg_cluster.cu), 4 and 16 CTAs
  dense     k_dense_solve, k_coarse_invert: per-pivot mbarriers instead of CTA barriers (dense_solve.cu, coarse.cu)
  coarse4   k_coarse_invert_cluster: four-CTA Gauss-Jordan, pivot rows stored into the peers' shared memory
  coarse16  k_coarse_invert_big: sixteen-CTA variant (MPBA_COARSE_GROUPS=48 set below, lagged setup on the side stream)
  lmgraph   the device LM loop: WHILE graph, k_lm_control / fused image tail, k_lm_accept
  integrate k_integrate: cluster CG with DSMEM halo rows and parity-double-buffered cluster sums (integrate.cu)
racecheck sees shared memory only (DSMEM accesses of peers are outside its model); memcheck covers global memory and
the cluster / mbarrier misuse checks.  GPU box only."""
import os
import sys

what = set(sys.argv[1:]) or {"all"}
if "coarse16" in what or "all" in what:
    os.environ["MPBA_COARSE_GROUPS"] = "48"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

from mpsfm_b200 import engine  # noqa: E402
from mpsfm_b200.synthetic import make_scene  # noqa: E402


def on(name):
    return "all" in what or name in what


def solve(tag, scene, **opt):
    ba = engine.BundleAdjuster(make_scene(*scene, seed=2))
    s = ba.solve(engine.default_options(**opt))
    print(f"{tag}: {s.brief_report()}", flush=True)
    ba.close()


if on("grid"):
    solve("grid PCG, block-Jacobi", (40, 600, 4), pcg_kernel=1, pcg_coarse=1, max_num_iterations=2, lm_loop=1)
    solve("grid PCG, two-level", (160, 1500, 4), pcg_kernel=1, max_num_iterations=2, lm_loop=1)
if on("cluster"):
    solve("cluster PCG, 4 CTAs", (40, 600, 4), pcg_kernel=2, max_num_iterations=2, lm_loop=1)
    solve("cluster PCG, 16 CTAs", (260, 2500, 4), pcg_kernel=2, max_num_iterations=2, lm_loop=1)
if on("dense"):
    solve("dense direct solve, 6 images", (6, 200, 3), max_num_iterations=2, lm_loop=1)
    solve("dense direct solve, 21 images", (21, 500, 4), max_num_iterations=2, lm_loop=1)
    solve("single-CTA coarse inverse (126 modes)", (300, 2500, 4), pcg_kernel=2, max_num_iterations=1, lm_loop=1)
if on("coarse4"):
    os.environ.pop("MPBA_COARSE_GROUPS", None)
    solve("four-CTA coarse inverse (224 modes), lagged", (640, 4000, 4), max_num_iterations=2, lm_loop=1)
if on("coarse16"):
    solve("sixteen-CTA coarse inverse (336 modes), grid PCG, lagged", (640, 4000, 4), pcg_kernel=1, max_num_iterations=2,
          lm_loop=1)
if on("lmgraph"):
    solve("device LM loop, dense", (8, 300, 4), lm_loop=2, max_num_iterations=4)
    solve("device LM loop, cluster PCG", (40, 600, 4), lm_loop=2, max_num_iterations=3)
if on("integrate"):
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden"))
    from make_golden_integration import make_inputs  # noqa: E402

    from mpsfm_b200 import integration as gi  # noqa: E402

    inp = make_inputs(3, 40, 56, 60)
    m = gi.MapInputs(depth_prior=inp["depth_prior"], uncertainty=inp["uncertainty"], valid=inp["valid"],
                     normals=inp["normals"], normals_uncertainty=inp["normals_uncertainty"], depth=inp["depth_init"],
                     K=inp["K"], kps=inp["kps"], depth3d=inp["depth3d"], zvars3d=inp["zvars3d"])
    r = gi.integrate_maps([m, m], gi.default_conf(max_iter=3, cg_max_iter=60))
    print("integrate:", r[0].updated, r[0].cg_iterations, bool(np.array_equal(r[0].depth, r[1].depth)), flush=True)
print("done")
