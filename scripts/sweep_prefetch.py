"""Sweep of the L2 lookahead prefetch distances of k_points / k_back_substitute / k_img_chunks (MPBA_POINTS_PREFETCH,
MPBA_BACKSUB_PREFETCH, MPBA_IMG_PREFETCH: CTAs ahead, 0 = off) on one workload; per-launch kernel times from the library's CUDA events.

    python scripts/sweep_prefetch.py [--workload cfg3] > profiles/r01_prefetch_sweep_cfg3.json
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from mpsfm_b200 import engine  # noqa: E402
from mpsfm_b200.synthetic import make_config  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="cfg3")
ap.add_argument("--iters", type=int, default=6)
ap.add_argument("--points", default="0,148,296,592,1184,2368,4736")
ap.add_argument("--backsub", default="0,333,666,1332,2664,5328")
ap.add_argument("--img", default="0,111,222,444,888")
a = ap.parse_args()
ba = engine.BundleAdjuster(make_config(a.workload, seed=0), device=0)


def run(pp, bp, ip=0):
    os.environ["MPBA_POINTS_PREFETCH"], os.environ["MPBA_BACKSUB_PREFETCH"] = str(pp), str(bp)
    os.environ["MPBA_IMG_PREFETCH"] = str(ip)
    best = None
    for _ in range(3):
        ba.reset_state()
        s = ba.solve(engine.default_options(max_num_iterations=a.iters, profile=1))
        t = {k: 1e3 * ms / n for k, (ms, n) in ba.kernel_times().items() if n}
        key = lambda d: d["points"] + d["back_substitute"] + d["img_chunks"]  # noqa: E731
        if best is None or key(t) < key(best):
            best = t
    return {"points_prefetch": pp, "backsub_prefetch": bp, "img_prefetch": ip, "points_us": round(best["points"], 1),
            "back_substitute_us": round(best["back_substitute"], 1), "schur_blocks_us": round(best["schur_blocks"], 1),
            "img_chunks_us": round(best["img_chunks"], 1), "final_cost": repr(s.final_cost)}


rows = [run(int(p), 0) for p in a.points.split(",")]
rows += [run(0, int(b)) for b in a.backsub.split(",")[1:]]
bp = min(rows[:len(a.points.split(","))], key=lambda r: r["points_us"])["points_prefetch"]
bb = min([rows[0]] + rows[len(a.points.split(",")):], key=lambda r: r["back_substitute_us"])["backsub_prefetch"]
rows += [run(0, 0, int(i)) for i in a.img.split(",")[1:]]
bi = min([rows[0]] + [r for r in rows if r["img_prefetch"]], key=lambda r: r["img_chunks_us"])["img_prefetch"]
rows.append(run(bp, bb, bi))
print(json.dumps({"workload": a.workload, "lm_iterations": a.iters, "rows": rows,
                  "bitwise_same_cost": len({r["final_cost"] for r in rows}) == 1}, indent=1))
