#!/usr/bin/env python
"""Golden vectors of the depth-integration path from the REFERENCE's own code.  Runs only where /root/reference exists.

/root/reference/mpsfm/sfm/scene/image/integration.py (class Integration), mpsfm/sfm/scene/camera.py (CameraIntData)
and mpsfm/utils/integration.py are imported UNMODIFIED, file by file (the package __init__ files pull in pycolmap).
The reference picks its matrix library at import time (mpsfm/utils/integration.py:6-26): cupy when
torch.cuda.is_available(), otherwise numpy + scipy.sparse + scipy.sparse.linalg.cg -- this container has no GPU, so
what runs here is the reference's own CPU mode.  Two imports it cannot satisfy are stubbed:
  cv2       only used by the `downscaled` branches (cv2.resize), which these cases do not take,
  cholespy  CholeskySolverF(n, ii, jj, data, CSR).solve(b, x): replaced by an exact scipy splu solve (the reference's
            variance propagation IntegrationUncertainty.solve stays the reference's code: one-hot right-hand sides,
            `x.sum(0)`).
The `cg` the reference calls is wrapped only to COUNT iterations (callback) and to snapshot z after every IRLS step.

    python tests/golden/make_golden_integration.py     # rewrites tests/golden/integration.npz

Each case stores the inputs (maps, sparse points, K, conf) and what Integration._integrate produced: return flag,
refined depth, cached wu / wv, energy_old, the energy of every IRLS step, CG iteration counts, z after each step; a
SECOND call on the updated state (the `integrated` branch: cached weights, energy-change gate); and for one case the
Hessian stencil + propagated variances (calculate_hessian with ignore_depths both ways, calculate_int_covs_at_points).
"""

from __future__ import annotations

import importlib.util
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"


def load_reference():
    import scipy.sparse as sp
    import scipy.sparse.linalg as spl
    import torch

    assert not torch.cuda.is_available(), "the golden vectors are defined by the reference's CPU (scipy) mode"
    cv2 = types.ModuleType("cv2")

    def _no_resize(*a, **k):
        raise RuntimeError("cv2.resize is not available: the golden cases do not take the downscaled branches")

    cv2.resize = _no_resize
    sys.modules["cv2"] = cv2
    ch = types.ModuleType("cholespy")

    class MatrixType:
        CSR = "CSR"

    class CholeskySolverF:
        def __init__(self, n, ii, jj, data, kind):
            A = sp.csr_matrix((data.cpu().numpy().astype(np.float64), jj.cpu().numpy(), ii.cpu().numpy()), shape=(n, n))
            self.lu = spl.splu(A.tocsc())

        def solve(self, b, x):
            x.copy_(torch.as_tensor(self.lu.solve(b.cpu().numpy().astype(np.float64)), dtype=x.dtype))

    ch.CholeskySolverF, ch.MatrixType = CholeskySolverF, MatrixType
    sys.modules["cholespy"] = ch
    for name in ("mpsfm", "mpsfm.sfm", "mpsfm.sfm.scene", "mpsfm.utils"):
        if name not in sys.modules:
            m = types.ModuleType(name)
            m.__path__ = []
            sys.modules[name] = m

    def load(name, rel):
        spec = importlib.util.spec_from_file_location(name, os.path.join(REF, rel))
        mod = importlib.util.module_from_spec(spec)
        sys.modules[name] = mod
        spec.loader.exec_module(mod)
        return mod

    load("mpsfm.utils.integration", "mpsfm/utils/integration.py")
    cam = load("mpsfm.sfm.scene.camera", "mpsfm/sfm/scene/camera.py")
    integ = load("mpsfm.sfm.scene.image.integration", "mpsfm/sfm/scene/image/integration.py")
    assert integ.device_g == "cpu"
    return integ, cam


class NS:
    def __init__(self, **kw):
        self.__dict__.update(kw)


DEFAULT_CONF = dict(large_number=1e6, max_iter=10, tol=5e-2, step_size=1, cg_max_iter=5000, cg_tol=1e-3, lambda1=1, lambda2=1,
                    k=1, depth_magnitude_multiplier=1, normals_magnitude_multiplier=1, cov_ignore_depth=True,
                    downscale_factor=2, downscaled=False, scale_filter=True, scale_filter_factor=1.5, robust_triangles=2,
                    ignore_depths=True, verbose=0)


def make_inputs(seed, H, W, n_sparse, noise=0.03, step=True, bias=0.06):
    """A smooth surface with a depth step, its analytic-ish normals (finite differences of the true depth, in the
    reference's (ny, nx, -nz) channel convention), a biased + noisy monocular depth prior, uncertainties, a validity
    mask with holes and sparse triangulated depths at integer pixels (with one duplicated pixel and one gross outlier
    for the scale filter)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    f = 1.2 * W
    K = [f, f, (H - 1) / 2.0 + 0.3, (W - 1) / 2.0 - 0.2]  # [K11 sy, K00 sx, K12 sy, K02 sx]
    r, c = np.meshgrid(np.arange(H), np.arange(W), indexing="ij")
    true = 4.0 + 0.6 * np.sin(c / W * 3.0) + 0.4 * np.cos(r / H * 2.5)
    if step:
        true = true + 1.2 * (c > 0.62 * W)
    logd = np.log(true)
    # normals from the log-depth gradients in the integration's own frame: u = (H-1-r) - cy (vertical), v = c - cx
    dzu = np.zeros((H, W))
    dzv = np.zeros((H, W))
    dzu[1:-1, :] = (logd[:-2, :] - logd[2:, :]) / 2.0   # d/du: u grows upwards (towards smaller r)
    dzv[:, 1:-1] = (logd[:, 2:] - logd[:, :-2]) / 2.0
    uu = (H - 1 - r) - K[2]
    vv = c - K[3]
    # n ~ (nx, ny, nz) with  dz/du = -nx / (uu nx + vv ny + f nz):  choose nz = -1 (facing the camera), solve 2x2
    nx = np.empty((H, W))
    ny = np.empty((H, W))
    for i in range(H):
        for j in range(W):
            a, b = dzu[i, j], dzv[i, j]
            M = np.array([[1 + a * uu[i, j], a * vv[i, j]], [b * uu[i, j], 1 + b * vv[i, j]]])
            rhs = np.array([a * f, b * f])
            nx[i, j], ny[i, j] = np.linalg.solve(M, rhs)
    nz = -np.ones((H, W))
    n = np.stack([nx, ny, nz], -1)
    n /= np.linalg.norm(n, axis=-1, keepdims=True)
    n += rng.normal(0, 0.01, n.shape)
    n /= np.linalg.norm(n, axis=-1, keepdims=True)
    normals = np.stack([n[..., 1], n[..., 0], -n[..., 2]], -1)  # channels the reference reads: [...,1]=nx, [...,0]=ny, -[...,2]=nz
    nunc = np.zeros((H, W, 3, 3))
    for k in range(3):
        nunc[..., k, k] = (0.02 + 0.02 * rng.random((H, W))) ** 2
    prior = true * (1.0 + bias * np.sin(c / W * 5.0 + seed)) * np.exp(rng.normal(0, noise, (H, W)))
    unc = (0.05 * prior) ** 2 * (0.5 + rng.random((H, W)))
    valid = rng.random((H, W)) > 0.04
    valid[H // 3: H // 3 + 3, W // 4: W // 4 + 4] = False
    xs = rng.integers(0, W, n_sparse)
    ys = rng.integers(0, H, n_sparse)
    if n_sparse >= 4:
        xs[1], ys[1] = xs[0], ys[0]  # a duplicated pixel: numpy's fancy += keeps the last one
    kps = np.stack([xs, ys], 1).astype(int)
    d3 = true[ys, xs] * np.exp(rng.normal(0, 0.01, n_sparse))
    if n_sparse >= 4:
        d3[2] *= 1.9  # removed by the scale filter
    zv = (0.01 * d3) ** 2 * (0.5 + rng.random(n_sparse))
    return dict(K=np.array(K), depth_prior=prior, uncertainty=unc, valid=valid, normals=normals, normals_uncertainty=nunc,
                depth_init=prior.copy(), kps=kps, depth3d=d3, zvars3d=zv)


def run_case(integ, cam, inp, conf_over, second_call_edit=None, with_hessian=False):
    conf = dict(DEFAULT_CONF)
    conf.update(conf_over)
    H, W = inp["depth_prior"].shape
    trace = {"cg": [], "z": []}
    real_cg = integ.cg

    def cg(A, b, x0=None, M=None, maxiter=None, rtol=1e-5):
        cnt = [0]

        def cb(xk):
            cnt[0] += 1

        x, info = real_cg(A, b, x0=x0, M=M, maxiter=maxiter, rtol=rtol, callback=cb)
        trace["cg"].append(cnt[0])
        trace["z"].append(np.array(x))
        return x, info

    integ.cg = cg
    energies = []
    real_energy = integ.Integration.calc_energy

    class Harness(integ.Integration):
        def calc_energy(self, *a, **k):
            e = real_energy(self, *a, **k)
            energies.append(float(e))
            return e

        def log(self, *a, **k):
            pass

    h = Harness()
    h.conf = NS(**conf)
    h.camera = cam.CameraIntData(H, W)
    h.depth = NS(data_prior=inp["depth_prior"].copy(), uncertainty=inp["uncertainty"].copy(), valid=inp["valid"].copy(),
                 data=inp["depth_init"].copy())
    h.normals = NS(data=inp["normals"].copy(), uncertainty=inp["normals_uncertainty"].copy())
    out = {}
    try:
        ok = h._integrate(inp["depth3d"].copy(), inp["zvars3d"].copy(), inp["kps"].copy(), list(inp["K"]))
        out.update(ok=bool(ok), depth=np.array(h.depth.data), wu=np.array(h.wu), wv=np.array(h.wv),
                   energy_old=float(h.energy_old), integrated=bool(h.integrated), energies=np.array(energies),
                   cg_iterations=np.array(trace["cg"]), z_steps=np.array(trace["z"]))
        # second call: the `integrated` branch
        if second_call_edit is not None:
            second_call_edit(h, inp)
            energies.clear()
            trace["cg"].clear()
            trace["z"].clear()
            depth_before = np.array(h.depth.data)
            ok2 = h._integrate(inp["depth3d"].copy(), inp["zvars3d"].copy(), inp["kps"].copy(), list(inp["K"]))
            out.update(second_ok=bool(ok2), second_depth_init=depth_before, second_depth=np.array(h.depth.data),
                       second_energy_old=float(h.energy_old), second_energies=np.array(energies),
                       second_cg_iterations=np.array(trace["cg"]), second_prior=np.array(h.depth.data_prior),
                       second_wu=np.array(h.wu), second_wv=np.array(h.wv))
        if with_hessian:
            integ.cg = real_cg
            h._prepare_integration_variables = lambda: (dict(kps=inp["kps"].copy(), zvars3d=inp["zvars3d"].copy(),
                                                             depth3d=inp["depth3d"].copy(), K=list(inp["K"])), True)
            qx = np.stack([np.arange(0, W, 3), (np.arange(0, W, 3) * 7) % H], 1).astype(float) + 0.2
            for tag, ign in (("ign", True), ("full", False)):
                h.Hessian = None
                h.calculate_hessian(downscaled=False, ignore_depths=ign)
                Hm = h.Hessian.tocsr()
                n = H * W
                idx = np.arange(n)
                out[f"hess_{tag}_diag"] = Hm.diagonal()
                er = np.zeros(n)
                ed = np.zeros(n)
                cmask = (idx % W) < W - 1
                er[cmask] = np.asarray(Hm[idx[cmask], idx[cmask] + 1]).reshape(-1)
                ed[: n - W] = np.asarray(Hm[idx[: n - W], idx[: n - W] + W]).reshape(-1)
                out[f"hess_{tag}_er"], out[f"hess_{tag}_ed"] = er, ed
                assert Hm.nnz <= 5 * n and abs(Hm - Hm.T).max() < 1e-9 * abs(Hm).max()
                out[f"var_{tag}"] = np.asarray(h.calculate_int_covs_at_points(qx, downscaled=False, ignore_depths=ign))
            out["var_query"] = qx
            out["hess_depth"] = np.array(h.depth.data)
    finally:
        integ.cg = real_cg
    return out


def main():
    integ, cam = load_reference()
    cases = {}

    def perturb_prior(h, inp):  # new evidence: the energy changes by more than tol -> refine again from cached weights
        h.depth.data_prior = h.depth.data_prior * 1.08

    def nothing(h, inp):  # nothing changed: "Energy hasn't changed. Skipping this frame."
        pass

    specs = [
        ("step_36x48", make_inputs(1, 36, 48, 40), {}, perturb_prior, True),
        ("smooth_30x40_skip", make_inputs(2, 30, 40, 25, step=False), {}, nothing, False),
        ("nosparse_24x32", make_inputs(3, 24, 32, 0), {"max_iter": 4}, None, False),
        ("tight_cg_40x56", make_inputs(4, 40, 56, 60, noise=0.05), {"cg_tol": 1e-8, "tol": 1e-3, "k": 2.0, "lambda2": 3.0,
                                                                    "normals_magnitude_multiplier": 2.0}, perturb_prior, False),
        ("noisy_energy_up_28x36", make_inputs(5, 28, 36, 30, noise=0.3, bias=0.4), {"max_iter": 6}, None, False),
    ]
    store = {}
    for name, inp, conf_over, second, hess in specs:
        out = run_case(integ, cam, inp, conf_over, second, hess)
        print(name, "ok", out["ok"], "energies", np.array2string(out["energies"], precision=4), "cg", out["cg_iterations"],
              "second", out.get("second_ok"), out.get("second_cg_iterations"))
        for k, v in inp.items():
            store[f"{name}/in/{k}"] = v
        for k, v in conf_over.items():
            store[f"{name}/conf/{k}"] = np.array(v)
        for k, v in out.items():
            store[f"{name}/out/{k}"] = np.asarray(v)
        cases[name] = 1
    store["cases"] = np.array(sorted(cases))
    np.savez_compressed(os.path.join(HERE, "integration.npz"), **store)
    print("wrote", os.path.join(HERE, "integration.npz"), os.path.getsize(os.path.join(HERE, "integration.npz")), "bytes")


if __name__ == "__main__":
    main()
