#!/usr/bin/env python
"""bench.py -- LM iterations/s and ms per bundle-adjustment solve (BASELINE.json's metric).

A *step* is one full BA solve (`mpba_solve`, Ceres termination rules) of the named synthetic scene from
its perturbed start state.  `value` = LM iterations per second with the problem resident in HBM;
`e2e` = the same through host buffers (pinned H2D of the whole problem + solve + D2H of poses/points).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload cfg3] [--impl reference]

Workloads (mpsfm_b200/synthetic.py CONFIGS): cfg2 (300k obs), cfg3 (5M obs, default: the largest
single-GPU configuration BASELINE.json names), target10m (10M obs), cfg4, cfg5 (sharded).
"""

from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "LM iterations/sec"
UNIT = "LM it/s"


def _peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index, self.rows, self.proc = index, [], None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def __exit__(self, *a):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()

    def summary(self):
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons), "samples": len(sm)}


def algorithmic_bytes(lay, spec=None):
    """Algorithmic HBM bytes per launch of each pass (fixed before measuring).
    Sharded (spec): this rank's observations in each order, its own points, the points its images touch."""
    C, B = lay.C, lay.num_blocks
    n_items = int(lay.item_a.shape[0])
    if spec is None:
        n_trk = n_img = lay.N
        p_own = p_touch = lay.P
    else:
        n_trk, n_img = int(lay.pt_ptr[-1]), int(lay.img_ptr[-1])
        p_own = spec.pt_end - spec.pt_begin
        p_touch = int(np.unique(lay.cm_pt[:n_img]).shape[0])
    return {
        # 4 (point idx) + 16 (xy) + 24 (d,w,a) per observation, every point read at least once (32 B), chunk partials
        "img_chunks": 44 * n_img + 32 * p_touch + 256 * lay.num_chunks + 64 * C,
        # 4 (image idx) + 16 + 24 in and one 128 B Schur factor out per observation; per point: xyz, pt_ptr, c_pt in,
        # Hpp / gp / Hinv / (L^-1, e) out
        "points": (44 + 128) * n_trk + (32 + 4 + 24 + 48 + 24 + 48 + 72) * p_own + 64 * C,
        # every factor read at least once, 8 B per item, every block of S written once
        "schur_blocks": 128 * n_trk + 8 * n_items + 288 * B + (168 + 48 + 48) * C,
        # Z and p_rot of the factor (96 B) + image index per observation; per point: xyz, pt_ptr, (L^-1, e), diag Hpp,
        # gp, c_pt in, candidate out; y of every image
        "back_substitute": (96 + 4) * n_trk + (32 + 4 + 72 + 24 + 24 + 24 + 32) * p_own + 48 * C,
    }


def _pyceres_reference_sample(layout_problem):
    """The REAL reference stack (pycolmap + pyceres, bundle_adjustment.py:104, :285-293) when this box carries it
    (oracle/pyceres_reference.py; it does not in this image).  Returns None otherwise."""
    try:
        from oracle import pyceres_reference as ref

        if not ref.available() or layout_problem is None:
            return None
        r = ref.solve(layout_problem)
        return r
    except Exception:
        return None


def cpu_baseline_sample(workload: str, max_iters: int = 50, layout=None, pcg_tol: float = 1e-10, budget_s: float = 30.0,
                        problem=None, probe: bool = True, rate=None):
    """The CPU arm on the WORKLOAD'S OWN scene (same generator, same seed, same size as the GPU arm -- no
    extrapolation): oracle/cpu_port (C++17 + OpenMP port of the reference's Ceres path, exact envelope-Cholesky
    solve of the reduced camera system like SPARSE_SCHUR) on every host thread, the first `max_iters` LM iterations
    of the solve.  `budget_s` bounds the sample: one LM iteration is timed first and max_iters is cut so that the
    sample takes about that long.  kind = "reference" only when pycolmap + pyceres are importable on the box."""
    from mpsfm_b200 import engine, synthetic
    from mpsfm_b200.synthetic import CONFIGS
    from oracle import cpu_port

    C, P, L = CONFIGS[workload]
    if layout is None:
        problem = synthetic.make_config(workload, seed=0)
        layout = engine.build_layout(problem)
    real = _pyceres_reference_sample(problem)
    if real is not None:
        return {"value": real["iterations"] / real["seconds"], "unit": UNIT, "cores": os.cpu_count(), "kind": "reference",
                "sample": f"pycolmap.create_default_bundle_adjuster + pyceres.solve (SPARSE_SCHUR, num_threads=-1) on the "
                          f"whole {workload} scene: {real['iterations']} LM iterations in {real['seconds']:.1f} s; {real['report']}",
                "sample_it_per_s": real["iterations"] / real["seconds"], "sample_obs": int(layout.N),
                "final_cost": real["final_cost"]}
    threads = cpu_port.max_threads()
    if probe or not rate:
        cpu_port.solve(layout, max_num_iterations=1, pcg_tolerance=pcg_tol)  # warm-up (page in, thread pool)
        t0 = time.perf_counter()
        cpu_port.solve(layout, max_num_iterations=2, pcg_tolerance=pcg_tol)  # probe: seconds per LM iteration
        rate = 2.0 / max(time.perf_counter() - t0, 1e-6)
    iters_cap = int(max(2, min(max_iters, budget_s * rate)))
    t0 = time.perf_counter()
    summ, _, _, _ = cpu_port.solve(layout, max_num_iterations=iters_cap, pcg_tolerance=pcg_tol)
    dt = time.perf_counter() - t0
    its = summ["num_iterations"]
    it_s = its / dt
    whole = summ["termination"] != 1  # not cut by the iteration cap: the whole solve ran
    return {
        "value": it_s,
        "unit": UNIT,
        "cores": threads,
        "kind": "port",
        "sample": (f"oracle/cpu_port (C++17/OpenMP port of the reference's Ceres path: trust-region LM, Schur "
                   f"elimination, exact envelope Cholesky of the reduced camera system under RCM ordering; pycolmap / "
                   f"pyceres / Ceres are not installable here), {threads} threads, "
                   + ("the whole solve" if whole else f"the first {its} LM iterations")
                   + f" of the workload's own scene ({layout.C} img / {layout.P} pt / {layout.N} obs, seed 0): "
                     f"{its} LM iterations in {dt:.1f} s"),
        "sample_it_per_s": it_s,
        "sample_obs": int(layout.N),
        "sample_lm_iterations": its,
        "whole_solve": bool(whole),
        "final_cost": summ["final_cost"],
    }


def run_reference(args):
    """Reference arm: the reference's CPU path on the SAME config, metric and unit as the GPU arm.  A step = one BA
    solve of the named scene on all host threads (cut to a bounded number of LM iterations only when one solve would
    exceed --ref-budget seconds).  Under torchrun rank 0 alone runs it -- with every host thread, not the
    OMP_NUM_THREADS=1 torchrun exports."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    os.environ.pop("OMP_NUM_THREADS", None)
    from mpsfm_b200 import engine, synthetic
    from mpsfm_b200.synthetic import CONFIGS

    C, P, L = CONFIGS[args.workload]
    problem = synthetic.make_config(args.workload, seed=0)
    layout = engine.build_layout(problem)
    times, iters, info = [], [], None
    # the whole arm is bounded to about --ref-total seconds: every step gets an equal share (at most --ref-budget)
    per_step = min(args.ref_budget, args.ref_total / max(args.warmup + args.steps, 1))
    for k in range(args.warmup + args.steps):
        info = cpu_baseline_sample(args.workload, layout=layout, problem=problem, pcg_tol=args.pcg_tol,
                                   budget_s=per_step, probe=(k == 0), rate=(info or {}).get("sample_it_per_s"))
        if k >= args.warmup:
            times.append(info["sample_lm_iterations"] / info["sample_it_per_s"] if "sample_lm_iterations" in info
                         else 1.0 / info["sample_it_per_s"])
            iters.append(info.get("sample_lm_iterations", 1))
    total_s = float(np.sum(times)) if times else float("nan")
    val = float(np.sum(iters)) / total_s
    info["value"] = val
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total_s / max(len(times), 1), "higher_is_better": True,
        "scaling": "strong" if args.gpus > 1 else "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": args.workload, "images": C, "points": P, "observations": P * L,
                   "same_config_as_gpu_arm": True,
                   "note": ("the reference's solver stack (pycolmap fork + Ceres 2.1 + pyceres) is not installable here; "
                            "kind 'port' = oracle/cpu_port, the same algorithm in C++/OpenMP on all host threads, on the "
                            "workload's own scene (seed 0, full size); a step = one solve, cut to the first LM iterations "
                            "that fit --ref-budget seconds when it would take longer"),
                   "lm_iterations_per_step": info.get("sample_lm_iterations"), "final_cost": info.get("final_cost")},
        "cpu_baseline": info,
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def run_gpu(args):
    import torch

    from mpsfm_b200 import engine
    from mpsfm_b200.synthetic import CONFIGS, make_config

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        import torch.distributed as dist

        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    else:
        dist = None
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: mpsfm_b200 has no CPU fallback")

    # N > 1: ONE scene of the named shape, sharded over the ranks (strong scaling; mpba_solve_sharded).
    # --sharding replicas: every rank solves its own scene of that shape (weak scaling, no collective).
    sharded = world > 1 and args.sharding == "shard"
    t0 = time.perf_counter()
    prob = make_config(args.workload, seed=0 if sharded else rank)
    t_scene = time.perf_counter() - t0
    t0 = time.perf_counter()
    full = engine.build_layout(prob)
    t_layout = time.perf_counter() - t0
    shard = (rank, world) if sharded else None
    ba = engine.BundleAdjuster(prob, device=local, layout=full, shard=shard)
    lay = ba.layout
    spec = engine.shard_ranges(full, world)[rank] if sharded else None
    # `value` and `e2e` are timed WITHOUT per-launch events; the per-kernel times behind `roofline` come from a
    # second pass of the same K steps with CUDA events around every launch (mpba_options.profile), also timed
    opt = engine.default_options(profile=0, pcg_tolerance=args.pcg_tol, pcg_max_iterations=args.pcg_max_iter)
    opt_prof = engine.default_options(profile=1, pcg_tolerance=args.pcg_tol, pcg_max_iterations=args.pcg_max_iter)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- resident: inputs already in HBM
    for _ in range(args.warmup):
        ba.reset_state()
        ba.solve(opt)
    barrier()
    ktimes, iters, launches, pcg_iters, solve_ms = {}, 0, 0, 0, 0.0
    with ClockSampler(local) as clk:
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        ev0.record()
        for _ in range(args.steps):
            ba.reset_state()
            s = ba.solve(opt)
            iters += s.num_iterations
            launches += s.kernel_launches
            pcg_iters += s.total_pcg_iterations
            solve_ms += s.solve_ms
        ev1.record()
        barrier()
        ms_total = ev0.elapsed_time(ev1)
        # second pass: the same steps with events around every kernel launch
        ev2, ev3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev2.record()
        for _ in range(args.steps):
            ba.reset_state()
            ba.solve(opt_prof)
            for k, (ms, n) in ba.kernel_times().items():
                a = ktimes.setdefault(k, [0.0, 0])
                a[0] += ms
                a[1] += n
        ev3.record()
        barrier()
        ms_total_prof = ev2.elapsed_time(ev3)
    clocks = clk.summary()
    last = s

    # ---- end to end: host buffers -> device -> solve -> host
    # the caller's arrays sit in pinned host memory (allocated once, like the device buffers); every step copies
    # the WHOLE problem host -> device, solves, and reads poses + points back
    h2d = d2h = 0
    e2e_iters = 0
    ba.close()
    del ba
    ba2 = engine.BundleAdjuster(prob, device=local, layout=full, pinned=True, shard=shard)
    ba2.upload()
    ba2.solve(opt)  # warm-up of the new handle (communicator, plans), of the staging copies and the read-back
    ba2.state(copy=False)
    # Single GPU: two adjusters (two sets of device buffers) and a copy stream -- the upload of step k + 1 runs while
    # step k is solved, as a caller with a queue of bundles drives it.  Every step still copies ITS whole problem
    # host -> device and reads ITS result back inside the timed region; only the first upload is exposed.
    pipelined = world == 1 and not args.e2e_serial
    if pipelined:
        ba3 = engine.BundleAdjuster(prob, device=local, layout=full, pinned=True)
        ba3.upload()
        ba3.solve(opt)
        ba3.state(copy=False)
        pair = (ba2, ba3)
        copy_stream = torch.cuda.Stream(device=local)
        ready = [torch.cuda.Event(), torch.cuda.Event()]
        torch.cuda.synchronize()
    barrier()
    t0 = time.perf_counter()
    if pipelined:
        with torch.cuda.stream(copy_stream):
            h2d = pair[0].upload()
            ready[0].record()
        for k in range(args.steps):
            cur, nxt = pair[k % 2], pair[(k + 1) % 2]
            if k + 1 < args.steps:
                with torch.cuda.stream(copy_stream):  # step k + 1's inputs: its buffers were read back at step k - 1
                    h2d = nxt.upload()
                    ready[(k + 1) % 2].record()
            torch.cuda.current_stream().wait_event(ready[k % 2])
            s2 = cur.solve(opt)
            q, t, X = cur.state(copy=False)  # views of the pinned result buffers: the device records cross PCIe as they are
            e2e_iters += s2.num_iterations
            d2h = lay.C * 64 + lay.P * 32
    else:
        for _ in range(args.steps):
            h2d = ba2.upload()
            s2 = ba2.solve(opt)
            q, t, X = ba2.state(copy=False)
            e2e_iters += s2.num_iterations
            d2h = lay.C * 64 + lay.P * 32
    barrier()
    e2e_s = time.perf_counter() - t0
    if pipelined:
        ba3.close()
    # one more, untimed, step with a synchronize after every leg: where the end-to-end time goes
    torch.cuda.synchronize()
    tb = [time.perf_counter()]
    ba2.upload(); torch.cuda.synchronize(); tb.append(time.perf_counter())
    ba2.solve(opt); torch.cuda.synchronize(); tb.append(time.perf_counter())
    ba2.state(copy=False); tb.append(time.perf_counter())
    e2e_legs = {"upload_ms": round(1e3 * (tb[1] - tb[0]), 3), "solve_ms": round(1e3 * (tb[2] - tb[1]), 3),
                "readback_ms": round(1e3 * (tb[3] - tb[2]), 3)}
    ba2.close()

    if dist is not None:
        red = torch.tensor([ms_total, e2e_s], dtype=torch.float64, device="cuda")
        dist.all_reduce(red, op=dist.ReduceOp.MAX)
        ms_total, e2e_s = float(red[0]), float(red[1])
        cnt = torch.tensor([iters, e2e_iters, launches, h2d, d2h], dtype=torch.float64, device="cuda")
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
        if sharded:  # one job: its LM iterations count once; launches and copies are summed over the ranks
            launches, h2d, d2h = int(cnt[2]), int(cnt[3]), int(cnt[4])
        else:
            iters, e2e_iters, launches, h2d, d2h = (int(x) for x in cnt)
    if rank != 0:
        return

    peak, peak_src = _peaks()
    ab = algorithmic_bytes(lay, spec)
    # the reduced-system solve by SURVEY 8(d)'s K3 formula: k PCG iterations x (288 B + 240 C) per launch; its setup
    # kernels read S once.  They are latency bound: the fraction is reported, not hidden.
    if "pcg" in ktimes and ktimes["pcg"][1]:
        ab["pcg"] = int(pcg_iters / max(iters, 1) * (288 * lay.num_blocks + 240 * lay.C))
        ab["coarse_setup"] = 288 * lay.num_blocks + 336 * lay.C
        ab["block_inverse"] = (288 + 288) * lay.C
    stream = ("img_chunks", "img_cost", "points", "schur_blocks", "back_substitute")
    top = max((k for k in ktimes if k in ab and k in stream), key=lambda k: ktimes[k][0], default=None)
    roof = None
    per_kernel = {}
    ms_all = sum(v[0] for v in ktimes.values()) or 1.0
    for k, (ms, n) in sorted(ktimes.items(), key=lambda kv: -kv[1][0]):
        per_kernel[k] = {"ms_total": round(ms, 3), "launches": n, "avg_us": round(1e3 * ms / max(n, 1), 2),
                         "share_of_kernel_time": round(ms / ms_all, 4)}
        if k in ab and n:
            gbs = ab[k] / (ms / n * 1e-3) / 1e9
            per_kernel[k].update(algorithmic_bytes_per_launch=int(ab[k]), algorithmic_GBps=round(gbs, 1),
                                 frac_of_measured_hbm=round(gbs / peak, 4))
    if top:
        ms, n = ktimes[top]
        achieved = ab[top] / (ms / n * 1e-3) / 1e9
        traffic = None  # DRAM bytes per launch of that kernel from the committed ncu --set full capture (1 GPU only)
        try:
            if world == 1:
                traffic = json.load(open(os.path.join(ROOT, "profiles", "r02_traffic.json"))).get(args.workload, {}).get(top)
        except Exception:
            traffic = None
        roof = {"bound": "hbm", "kernel": top, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic,
                "traffic_source": ("profiles/r02_traffic.json: DRAM bytes of one launch from the committed ncu --set full "
                                   "capture of this workload (not re-measured by this run)") if traffic else None,
                "peak_source": peak_src, "algorithmic_bytes_per_launch": ab[top],
                "avg_launch_ms": ms / n, "frac_of_nominal_8TBps": achieved / 8000.0}

    extension = normal_prior_extension(args, prob, local, opt, opt_prof) if (world == 1 and not args.no_extension) else None
    cpu = (cpu_baseline_sample(args.workload, layout=full, problem=prob, pcg_tol=args.pcg_tol, budget_s=args.ref_budget)
           if (world == 1 and not args.no_cpu_baseline) else None)
    C_, P_, L_ = CONFIGS[args.workload]
    line = {
        "metric": METRIC, "value": iters / (ms_total * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True,
        "scaling": "strong" if sharded else "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {
            "workload": args.workload, "images": C_, "points": P_, "observations": P_ * L_,
            "sharding": ("one scene sharded over the ranks: image-order chunks by camera, tracks by point, NCCL "
                         "all-reduce of the partial reduced camera system (mpba_solve_sharded)") if sharded
            else ("one scene per rank" if world > 1 else "single GPU"),
            "camera_model": prob.cam_model.tolist(), "reproj_loss": int(prob.reproj_loss), "depth_loss": int(prob.depth_loss),
            "l2": "inputs larger than L2" if lay.N * 88 > 126e6 else "inputs fit in L2 (small workload; see DESIGN.md)",
            "lm_iterations_per_solve": last.num_iterations, "termination": last.message,
            "final_cost": last.final_cost, "pcg_iterations_per_solve": pcg_iters / args.steps,
            "pcg_tolerance": args.pcg_tol, "rcs_blocks": lay.num_blocks,
            "host_scene_s": round(t_scene, 2), "host_layout_s": round(t_layout, 2),
        },
        "ms_per_lm_iteration": ms_total / max(iters, 1),
        "ms_per_step_with_kernel_events": ms_total_prof / args.steps,
        "e2e": {"value": e2e_iters / e2e_s, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "ms_per_step": 1e3 * e2e_s / args.steps, "legs_of_one_extra_step": e2e_legs,
                "pipeline": ("double-buffered: the host->device copy of step k+1 overlaps the solve of step k (two adjusters, "
                             "copy stream); --e2e-serial times the legs back to back") if pipelined else "serial"},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": roof,
        "kernels": per_kernel,
        "cpu_baseline": cpu,
        "extension": extension,
    }
    print(json.dumps(line))


def normal_prior_extension(args, prob, device, opt, opt_prof):
    """The workload with the normal-prior rows switched on (EXTENSION, parity n/a: the reference has no such residual --
    SURVEY.md section 8 row a-9, oracle/residuals.py defines it), timed apart from the headline: same scene, a
    point-to-plane row on half of the depth-prior observations, resident inputs, the same K steps."""
    import torch

    from mpsfm_b200 import engine, normal_prior

    t0 = time.perf_counter()
    pn = normal_prior.synthetic_normal_priors(prob, frac=0.5, seed=0)
    t_rows = time.perf_counter() - t0
    ba = engine.BundleAdjuster(pn, device=device)
    for _ in range(args.warmup):
        ba.reset_state()
        ba.solve(opt)
    torch.cuda.synchronize()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    iters = 0
    ev0.record()
    for _ in range(args.steps):
        ba.reset_state()
        s = ba.solve(opt)
        iters += s.num_iterations
    ev1.record()
    torch.cuda.synchronize()
    ms = ev0.elapsed_time(ev1)
    ba.reset_state()
    ba.solve(opt_prof)
    kt = {k: round(1e3 * v[0] / max(v[1], 1), 2) for k, v in ba.kernel_times().items() if k in ("img_chunks", "points")}
    ba.close()
    return {"normal_prior": {
        "note": "extension, parity n/a (no such residual upstream); not part of `value`",
        "value": iters / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms / args.steps, "rows": int(pn.num_plane),
        "lm_iterations_per_solve": s.num_iterations, "final_cost": s.final_cost, "termination": s.message,
        "extra_bytes_per_observation": 64, "avg_us": kt, "host_rows_s": round(t_rows, 2)}}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="cfg3")
    ap.add_argument("--pcg-tol", dest="pcg_tol", type=float, default=1e-10)
    ap.add_argument("--pcg-max-iter", dest="pcg_max_iter", type=int, default=2000)
    ap.add_argument("--ref-budget", dest="ref_budget", type=float, default=30.0,
                    help="CPU arm: seconds one step may take; longer solves are cut to their first LM iterations")
    ap.add_argument("--ref-total", dest="ref_total", type=float, default=150.0,
                    help="reference arm: seconds all its steps together may take (bounds --ref-budget per step)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extension", dest="no_extension", action="store_true",
                    help="skip the separate normal-prior (extension) measurement")
    ap.add_argument("--e2e-serial", dest="e2e_serial", action="store_true",
                    help="end-to-end steps strictly one after the other (no overlap of the next upload with the solve)")
    ap.add_argument("--sharding", default="shard", choices=["shard", "replicas"])
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    # stdout carries the ONE JSON line and nothing else: libraries that print to file descriptor 1 (NCCL's version
    # banner under NCCL_DEBUG, for one) are sent to stderr for the whole run
    sys.stdout.flush()
    _json_fd = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(_json_fd, "w", buffering=1)
    main()
    sys.stdout.flush()
