"""Multi-threaded CPU port of the bundle-adjustment path (C++17 + OpenMP) -- TIMING BASELINE, not a parity oracle.

TEST / BENCHMARK INFRASTRUCTURE ONLY: ``bench.py`` (cpu_baseline, ``--impl reference``) and ``tests/`` may import
this; nothing under ``mpsfm_b200/`` does.  See ba_cpu.cpp for what it restates and how it differs from Ceres.
The problem is handed over in the array layout of ``include/mpba.h`` with HOST pointers."""

from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "ba_cpu.cpp")
LIB = os.path.join(HERE, "libba_cpu.so")
_DEPS = (SRC, os.path.join(HERE, "..", "..", "mpsfm_b200", "csrc", "ba_math.cuh"), os.path.join(HERE, "..", "..", "include", "mpba.h"))
_lib = None


def build(force: bool = False) -> str:
    """g++ -O3 -fopenmp.  Built -march=x86-64-v3 (AVX2 + FMA) HERE, because the library travels to the GPU box
    prebuilt and that box's host CPU is a different model (-march=native of the build container could emit
    instructions it lacks); set MPBA_CPU_PORT_NATIVE=1 to build -march=native on the machine that will run it
    (bench.py's reference arm does, and falls back to the prebuilt library when no compiler is there)."""
    native = os.environ.get("MPBA_CPU_PORT_NATIVE") == "1"
    lib = LIB.replace(".so", "_native.so") if native else LIB
    if force or not os.path.exists(lib) or any(os.path.getmtime(d) > os.path.getmtime(lib) for d in _DEPS):
        subprocess.check_call(["g++", "-O3", "-march=native" if native else "-march=x86-64-v3", "-fopenmp", "-std=c++17",
                               "-fPIC", "-shared", "-x", "c++", SRC, "-o", lib])
    return lib


def _load():
    global _lib
    if _lib is None:
        from mpsfm_b200 import engine

        try:
            path = build()
        except (OSError, subprocess.CalledProcessError):
            if not os.path.exists(LIB):
                raise
            path = LIB  # no compiler on this box: the prebuilt x86-64-v3 library
        _lib = C.CDLL(path)
        _lib.ba_cpu_solve.argtypes = [C.POINTER(engine.MpbaProblem), C.POINTER(engine.MpbaOptions),
                                      C.POINTER(engine.MpbaSummary), C.c_int]
        _lib.ba_cpu_max_threads.restype = C.c_int
    return _lib


def max_threads() -> int:
    """Host threads the port runs on by default: every CPU this process may use.  (NOT omp_get_max_threads():
    torchrun exports OMP_NUM_THREADS=1 to its workers, which would time the reference arm on one core.)"""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return os.cpu_count() or 1


def solve(layout, max_num_iterations: int = 50, pcg_tolerance: float = 1e-10, num_threads: int = 0,
          linear_solver: str = "cholesky"):
    """Run the LM loop on a ``mpsfm_b200.engine.HostLayout``.  Returns (summary dict, q [C,4], t [C,3], X [P,3]);
    the layout's arrays are not modified.  ``linear_solver``: "cholesky" (exact envelope Cholesky of the reduced
    camera system, what Ceres' SPARSE_SCHUR does) or "pcg" (block-Jacobi PCG to ``pcg_tolerance``).
    ``num_threads`` 0 = every hardware thread (torchrun exports OMP_NUM_THREADS=1: that is overridden here)."""
    from mpsfm_b200 import engine

    lib = _load()
    lib.ba_cpu_set_linear_solver(1 if linear_solver == "pcg" else 0)
    if num_threads <= 0:
        num_threads = max_threads()
    keep = {}
    prob = engine.MpbaProblem()
    for name in engine._DEVICE_FIELDS:
        arr = np.ascontiguousarray(getattr(layout, name))
        if name in ("img_pose", "pt_xyz"):
            arr = arr.copy()
        if arr.size == 0:
            arr = np.zeros(1, dtype=arr.dtype)
        keep[name] = arr
        setattr(prob, name, arr.ctypes.data)
    prob.num_images, prob.num_cameras, prob.num_points, prob.num_obs = layout.C, layout.K, layout.P, layout.N
    prob.num_chunks, prob.num_blocks = layout.num_chunks, layout.num_blocks
    prob.num_upper = int(layout.u_block.shape[0])
    prob.num_track_warps = int(layout.track_warp_first.shape[0]) - 1
    prob.reproj_loss, prob.depth_loss = layout.reproj_loss, layout.depth_loss
    prob.reproj_loss_scale, prob.reproj_loss_weight = layout.reproj_loss_scale, layout.reproj_loss_weight
    prob.flags = layout.flags
    if getattr(layout, "pm_plane", None) is not None:  # normal-prior rows (extension)
        keep["pm_plane"], keep["cm_plane"] = np.ascontiguousarray(layout.pm_plane), np.ascontiguousarray(layout.cm_plane)
        prob.pm_plane, prob.cm_plane = keep["pm_plane"].ctypes.data, keep["cm_plane"].ctypes.data
        prob.plane_loss_scale = layout.plane_loss_scale
    opt = engine.MpbaOptions(
        max_num_iterations=max_num_iterations, max_num_consecutive_invalid_steps=5, jacobi_scaling=1,
        pcg_max_iterations=50000, initial_trust_region_radius=1e4, max_trust_region_radius=1e16,
        min_trust_region_radius=1e-32, min_relative_decrease=1e-3, min_lm_diagonal=1e-6, max_lm_diagonal=1e32,
        function_tolerance=1e-6, gradient_tolerance=1e-10, parameter_tolerance=1e-8, pcg_tolerance=pcg_tolerance)
    cap = max_num_iterations + 2
    rows = (engine.MpbaIteration * cap)()
    sm = engine.MpbaSummary()
    sm.iterations = C.cast(rows, C.POINTER(engine.MpbaIteration))
    sm.capacity = cap
    rc = lib.ba_cpu_solve(C.byref(prob), C.byref(opt), C.byref(sm), int(num_threads))
    if rc != 0:
        raise RuntimeError(f"ba_cpu_solve failed with code {rc}")
    n = min(sm.num_iterations + 1, cap)
    summary = {
        "num_iterations": sm.num_iterations, "initial_cost": sm.initial_cost, "final_cost": sm.final_cost,
        "termination": sm.termination, "message": sm.message.decode(), "total_pcg_iterations": int(sm.total_pcg_iterations),
        "costs": [rows[k].cost for k in range(n)], "successful": [rows[k].step_is_successful for k in range(n)],
    }
    pose, xyz = keep["img_pose"], keep["pt_xyz"]
    return summary, pose[:, :4].copy(), pose[:, 4:7].copy(), xyz[:, :3].copy()
