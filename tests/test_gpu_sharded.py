"""Sharded solve (include/mpba.h: mpba_solve_sharded) against the single-GPU solve of the same scene.

Gate (SURVEY.md section 8 row e, "Determinism"): the N-GPU trajectory agrees with the 1-GPU one to fp64
round-off -- same LM iteration count and accept / reject sequence, costs <= 1e-10 relative, poses / points
<= 1e-8 -- and every rank ends with the same solution bit for bit."""

import json
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

from mpsfm_b200 import engine
from mpsfm_b200.problem import CameraModel, LossType
from mpsfm_b200.synthetic import make_scene

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("pcg_kernel", [0, 1])
def test_world_of_one_runs_the_sharded_code_path(pcg_kernel):
    """world = 1: partial Schur blocks + finish kernel + NCCL collectives (trivial) on one GPU."""
    prob = make_scene(16, 1200, 5, seed=2, model=CameraModel.SIMPLE_RADIAL, reproj_loss=LossType.CAUCHY,
                      const_point_frac=0.05)
    lay = engine.build_layout(prob)
    opt = engine.default_options(pcg_kernel=pcg_kernel)
    a = engine.BundleAdjuster(prob, device=0, layout=lay)
    sa = a.solve(opt)
    b = engine.BundleAdjuster(prob, device=0, layout=lay, shard=(0, 1))
    sb = b.solve(opt)
    assert sa.num_iterations == sb.num_iterations
    ca = np.array([it["cost"] for it in sa.iterations])
    cb = np.array([it["cost"] for it in sb.iterations])
    assert np.max(np.abs(ca - cb) / ca) <= 1e-12
    for x, y in zip(a.state(), b.state()):
        assert np.max(np.abs(x - y)) <= 1e-9
    with pytest.raises(engine.EngineError):
        b.evaluate()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("extra", [[], ["--images", "40", "--points", "6000", "--pcg-kernel", "1"]])
def test_two_gpus_match_one(extra):
    import torch

    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs two GPUs (gpurun --gpus 2)")
    world = 2
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()),
           os.path.join(ROOT, "tests", "sharded_worker.py"), *extra]
    out = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=600)
    lines = [ln for ln in out.stdout.splitlines() if ln.startswith("{")]
    assert out.returncode == 0 and lines, out.stdout[-2000:] + out.stderr[-2000:]
    res = json.loads(lines[-1])
    assert res["ok"] and res["replicas_identical"], res


def test_structure_only_through_the_sharded_path():
    """refine_3d_points (all poses constant): no reduced system, only the image-block all-reduce and the point
    all-gather remain."""
    prob = make_scene(10, 500, 5, seed=2, depth_loss=LossType.TRIVIAL)
    prob.img_free_mask[:] = 0
    lay = engine.build_layout(prob)
    a = engine.BundleAdjuster(prob, device=0, layout=lay)
    sa = a.solve()
    b = engine.BundleAdjuster(prob, device=0, layout=lay, shard=(0, 1))
    sb = b.solve()
    assert sa.num_iterations == sb.num_iterations and abs(sa.final_cost - sb.final_cost) <= 1e-12 * sa.final_cost
    assert np.max(np.abs(a.state()[2] - b.state()[2])) <= 1e-10
