"""Latency of small bundles (the reference's own example is 6 images / ~10^4 observations): host layout,
buffer creation, solve and read-back, next to the CPU port on the same problem.

    python scripts/bench_small.py > profiles/r02_small_bundles.json

Round 2: the solve is timed with the LM controller on the host (lm_loop = 1, round 1's path) and on the device
(lm_loop = 0 picks it below 600 images: one CUDA-graph launch per solve), the latter on a fresh adjuster (the graph
is captured and instantiated inside the timed call) and on a kept one (graph reused).
"""

import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mpsfm_b200 import engine  # noqa: E402
from mpsfm_b200.synthetic import make_scene  # noqa: E402
from oracle import cpu_port  # noqa: E402


def main():
    out = []
    for C, P, L in ((6, 3000, 4), (20, 10000, 5), (50, 30000, 5), (100, 50000, 6)):
        prob = make_scene(C, P, L, seed=0)
        t0 = time.perf_counter()
        lay = engine.build_layout(prob)
        t_layout = time.perf_counter() - t0
        engine.BundleAdjuster(prob, layout=lay).solve()  # warm-up (context, plans)
        reps = 5
        t_make = t_read = 0.0
        t_solve = {"host": 0.0, "device_fresh": 0.0, "device_kept": 0.0}
        dev_ms = {"host": 0.0, "device_fresh": 0.0, "device_kept": 0.0}
        for _ in range(reps):
            for mode, loop in (("host", 1), ("device_fresh", 0)):
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                ba = engine.BundleAdjuster(prob, layout=lay)
                torch.cuda.synchronize()
                t1 = time.perf_counter()
                s = ba.solve(engine.default_options(lm_loop=loop))
                t2 = time.perf_counter()
                ba.state()
                t3 = time.perf_counter()
                t_make += (t1 - t0) / 2
                t_solve[mode] += t2 - t1
                dev_ms[mode] += s.solve_ms
                t_read += (t3 - t2) / 2
                if mode == "device_fresh":
                    ba.reset_state()
                    torch.cuda.synchronize()
                    t2 = time.perf_counter()
                    s = ba.solve(engine.default_options(lm_loop=loop))
                    t_solve["device_kept"] += time.perf_counter() - t2
                    dev_ms["device_kept"] += s.solve_ms
                ba.close()
        cpu_port.solve(lay, num_threads=0)
        t0 = time.perf_counter()
        cs, _, _, _ = cpu_port.solve(lay, num_threads=0)
        t_cpu = time.perf_counter() - t0
        out.append({"images": C, "points": P, "observations": int(lay.N), "lm_iterations": s.num_iterations,
                    "host_layout_ms": 1e3 * t_layout, "create_upload_ms": 1e3 * t_make / reps,
                    "solve_ms": {k: 1e3 * v / reps for k, v in t_solve.items()},
                    "solve_ms_cuda_events": {k: v / reps for k, v in dev_ms.items()},
                    "us_per_lm_iteration": {k: 1e6 * v / reps / max(s.num_iterations, 1) for k, v in t_solve.items()},
                    "readback_ms": 1e3 * t_read / reps, "kernel_launches": s.kernel_launches,
                    "cpu_port_ms": 1e3 * t_cpu, "cpu_port_iterations": cs["num_iterations"], "cpu_threads": cpu_port.max_threads()})
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
