"""Depth-integration solver (mpba_integrate_depth, csrc/integrate.cu) next to the CPU restatement of the reference's
solver (oracle/integration.py: numpy stencils + the scipy-semantics CG the reference runs through cupy, host cores).
Reference resolution: long side 387 px (mpsfm/sfm/scene/reconstruction/base.py:27).  GPU box only.

    python scripts/bench_integration.py > profiles/r02_integration.json
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import numpy as np  # noqa: E402
import torch  # noqa: E402
from make_golden_integration import make_inputs  # noqa: E402

from mpsfm_b200 import integration as gi  # noqa: E402
from oracle import integration as oi  # noqa: E402


def maps_of(inp):
    return gi.MapInputs(depth_prior=inp["depth_prior"], uncertainty=inp["uncertainty"], valid=inp["valid"],
                        normals=inp["normals"], normals_uncertainty=inp["normals_uncertainty"], depth=inp["depth_init"],
                        K=inp["K"], kps=inp["kps"], depth3d=inp["depth3d"], zvars3d=inp["zvars3d"])


def main():
    H, W = 290, 387
    inputs = [make_inputs(100 + k, H, W, 2500) for k in range(8)]
    conf = gi.default_conf()
    gi.integrate_maps([maps_of(inputs[0])], conf)  # warm-up
    out = {"grid": [H, W], "unknowns_per_image": H * W, "rows": []}
    for batch in (1, 8, 32):
        maps = [maps_of(inputs[k % len(inputs)]) for k in range(batch)]
        best = 1e9
        for _ in range(3):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            res = gi.integrate_maps(maps, conf)
            best = min(best, time.perf_counter() - t0)
        cg = int(sum(sum(r.cg_iterations) for r in res))
        out["rows"].append({"images": batch, "ms_per_call_including_copies": 1e3 * best, "ms_per_image": 1e3 * best / batch,
                            "cg_iterations_total": cg, "irls_steps": [len(r.cg_iterations) for r in res][:4]})
    oconf = oi.IntegrationConf()
    inp = inputs[0]
    t0 = time.perf_counter()
    trace = {}
    pr = oi.prepare(inp["depth_prior"], inp["uncertainty"], inp["valid"], inp["normals"], inp["normals_uncertainty"],
                    inp["depth_init"], inp["kps"], inp["depth3d"], inp["zvars3d"], inp["K"], oconf)
    ok, depth, _ = oi.integrate(pr, oconf, trace=trace)
    t_cpu = time.perf_counter() - t0
    g = gi.integrate_maps([maps_of(inp)], conf)[0]
    out["cpu_oracle"] = {"ms_per_image": 1e3 * t_cpu, "cg_iterations_total": int(sum(trace["cg_iterations"])),
                         "what": "oracle/integration.py (numpy 5-point stencils, CG with scipy's stopping rule), one image"}
    out["same_result"] = {"cg_iterations_equal": list(g.cg_iterations) == list(trace["cg_iterations"]),
                          "max_rel_log_depth_diff": float(np.max(np.abs(np.log(g.depth) - np.log(depth))) /
                                                          np.max(np.abs(np.log(depth))))}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
