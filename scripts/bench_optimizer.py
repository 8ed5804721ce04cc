"""Whole ``Optimizer.ba`` call on the pycolmap-shaped fake scene of the tests (tests/scene_factory.py): assembly, host
layout, buffers + upload, solve, read-back and the in-place write-back, timed separately and as one call.  GPU box only.

    python scripts/bench_optimizer.py > profiles/r02_optimizer_call.json
"""
import copy
import cProfile
import json
import os
import pstats
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402
from scene_factory import build_scene  # noqa: E402

from mpsfm_b200 import engine  # noqa: E402
from mpsfm_b200.optimizer import Optimizer  # noqa: E402


def main():
    out = []
    for C, P in ((6, 3000), (20, 10000), (50, 30000)):
        rec0 = build_scene(seed=1, num_images=C, num_points=P)

        def fresh():  # every timed call starts from the same perturbed scene (ba updates it in place)
            rec = copy.deepcopy(rec0)
            opt = Optimizer({}, rec, None)
            opt.update_truncation_multiplier(rec.reg_image_ids())
            return rec, opt, {"optim_ids": list(rec.reg_image_ids()), "pts3D": set(rec.points3D)}

        rec, opt, bundle = fresh()
        opt.ba(bundle, mode="global", allow_scale_filter=True)  # warm-up: library, context, kernels
        reps = 3
        whole = 0.0
        for _ in range(reps):
            rec, opt, bundle = fresh()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            opt.ba(bundle, mode="global", allow_scale_filter=True)
            whole += time.perf_counter() - t0
        rec, opt, bundle = fresh()
        if "--profile" in sys.argv and C == 20:
            pr = cProfile.Profile()
            pr.enable()
            opt.assemble(bundle, fix_pose=False, mode="global", allow_scale_filter=True)
            pr.disable()
            pstats.Stats(pr, stream=sys.stderr).sort_stats("tottime").print_stats(18)
        # the same call taken apart
        t0 = time.perf_counter()
        asm = opt.assemble(bundle, fix_pose=False, mode="global", allow_scale_filter=True)
        t1 = time.perf_counter()
        lay = engine.build_layout(asm.problem)
        t2 = time.perf_counter()
        ba = engine.BundleAdjuster(asm.problem, layout=lay)
        torch.cuda.synchronize()
        t3 = time.perf_counter()
        s = ba.solve()
        t4 = time.perf_counter()
        q, t, X = ba.state()
        t5 = time.perf_counter()
        opt._write_back(asm, q, t, X)
        t6 = time.perf_counter()
        ba.close()
        out.append({"images": C, "points": P, "observations": int(asm.problem.obs_img.shape[0]),
                    "lm_iterations": s.num_iterations, "optimizer_ba_ms": 1e3 * whole / reps,
                    "parts_ms": {"assemble": 1e3 * (t1 - t0), "host_layout": 1e3 * (t2 - t1),
                                 "buffers_upload": 1e3 * (t3 - t2), "solve": 1e3 * (t4 - t3),
                                 "readback": 1e3 * (t5 - t4), "write_back": 1e3 * (t6 - t5)}})
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
