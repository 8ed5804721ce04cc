"""Worker of tests/test_gpu_sharded.py, one process per GPU:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        tests/sharded_worker.py [--images C --points P --track L --model M]

Every rank solves the whole scene alone on its GPU (mpba_solve) and then its shard together with the others
(mpba_solve_sharded); the two trajectories must agree to fp64 round-off and every rank must end with the same
full solution.  Prints one JSON line on rank 0; exit code 1 on a mismatch."""

import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist

    from mpsfm_b200 import engine
    from mpsfm_b200.problem import CameraModel, LossType
    from mpsfm_b200.synthetic import make_scene

    ap = argparse.ArgumentParser()
    ap.add_argument("--images", type=int, default=24)
    ap.add_argument("--points", type=int, default=3000)
    ap.add_argument("--track", type=int, default=5)
    ap.add_argument("--model", default="SIMPLE_RADIAL")
    ap.add_argument("--pcg-kernel", type=int, default=0)
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    prob = make_scene(args.images, args.points, args.track, seed=5, model=getattr(CameraModel, args.model),
                      reproj_loss=LossType.CAUCHY, const_point_frac=0.05, depth_frac=0.8)
    lay = engine.build_layout(prob)
    opt = engine.default_options(pcg_kernel=args.pcg_kernel)
    one = engine.BundleAdjuster(prob, device=local, layout=lay)
    s1 = one.solve(opt)
    q1, t1, X1 = one.state()
    sh = engine.BundleAdjuster(prob, device=local, layout=lay, shard=(rank, world))
    sN = sh.solve(opt)
    qN, tN, XN = sh.state()

    c1 = np.array([it["cost"] for it in s1.iterations])
    cN = np.array([it["cost"] for it in sN.iterations])
    res = {
        "world": world, "rank": rank,
        "iterations": [s1.num_iterations, sN.num_iterations],
        "accept": [[it["step_is_successful"] for it in s.iterations] for s in (s1, sN)],
        "cost_rel": float(np.max(np.abs(c1 - cN) / c1)) if c1.shape == cN.shape else None,
        "final_rel": abs(s1.final_cost - sN.final_cost) / s1.final_cost,
        "dq": float(np.max(np.abs(q1 - qN))), "dt": float(np.max(np.abs(t1 - tN))), "dX": float(np.max(np.abs(X1 - XN))),
        "launches": sN.kernel_launches,
    }
    # every rank holds the same full solution, bit for bit (replicated solve of the reduced system)
    mine = torch.from_numpy(np.concatenate([qN.ravel(), tN.ravel(), XN.ravel()])).cuda()
    ref = mine.clone()
    dist.broadcast(ref, src=0)
    res["replicas_identical"] = bool(torch.equal(mine, ref))
    ok = (res["iterations"][0] == res["iterations"][1] and res["accept"][0] == res["accept"][1]
          and res["cost_rel"] is not None and res["cost_rel"] <= 1e-10 and res["final_rel"] <= 1e-10
          and res["dq"] <= 1e-8 and res["dt"] <= 1e-8 and res["dX"] <= 1e-7 and res["replicas_identical"])
    # one rank fails its LOCAL checks (workspace too small): every rank must come back with an error instead of the
    # others waiting in their first collective, and the communicator must still be usable afterwards
    keep = sh.workspace
    if rank == world - 1:
        sh.workspace = torch.empty(256, dtype=torch.uint8, device="cuda")
    sh.reset_state()
    try:
        sh.solve(opt)
        res["local_failure_reported"] = False
    except engine.EngineError as e:
        res["local_failure_reported"] = ("workspace too small" in str(e)) if rank == world - 1 else ("another rank" in str(e))
    sh.workspace = keep
    sh.reset_state()
    again = sh.solve(opt)
    res["solve_after_failure"] = again.num_iterations == sN.num_iterations and again.final_cost == sN.final_cost
    ok = ok and res["local_failure_reported"] and res["solve_after_failure"]
    flags = torch.tensor([1.0 if ok else 0.0], device="cuda")
    dist.all_reduce(flags, op=dist.ReduceOp.MIN)
    if rank == 0:
        res["ok"] = bool(flags.item() == 1.0)
        del res["accept"]
        print(json.dumps(res), flush=True)
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if flags.item() == 1.0 else 1)


if __name__ == "__main__":
    main()
