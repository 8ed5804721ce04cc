"""Golden vectors of the numpy oracle on the NAMED BASELINE configs.  TEST INFRASTRUCTURE.

The oracle needs ~2 minutes for a cfg2 solve and ~10 minutes / ~20 GB for one cfg3 linearisation, too long for the GPU
box's test budget, so its outputs are stored here once and the `-m gpu` tests compare the CUDA path with them:

  cfg2_trajectory.npz        oracle/lm.solve on cfg2 (100 images / 50k points / 300k observations, PINHOLE, SoftL1 +
                             Cauchy): per-iteration cost, accept / reject flags, radius, final poses and a point sample
  cfg3_first_iteration.npz   cfg3 (1000 images / 1M points / 5M observations, SIMPLE_RADIAL, Cauchy): residual sample,
                             cost, Hcc / gc of every image, Hpp / gp sample, a 48 x 48 sample of the reduced camera
                             system and the full rhs at radius 1e4, the first LM step (y_c, y_p sample), the candidate
                             cost and the accept decision

Run from the repo root:  python tests/golden/make_golden_configs.py [cfg2] [cfg3]
The scenes come from the seeded generator (mpsfm_b200/synthetic.py: make_config(name, seed=0)); the oracle is
deterministic numpy, so the files are reproducible.  These are ORACLE outputs: like the oracle they are unpinned
against Ceres itself.
"""

from __future__ import annotations

import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from mpsfm_b200.synthetic import make_config  # noqa: E402
from oracle import lm  # noqa: E402
from oracle.residuals import State, evaluate, free_columns  # noqa: E402

S_SAMPLE_IMAGES = np.array([1, 2, 3, 250, 251, 500, 998, 999])


def cfg2():
    prob = make_config("cfg2")
    t = time.time()
    s = lm.solve(prob)
    print(f"cfg2: {len(s.iterations) - 1} iterations, final cost {s.final_cost!r}, {time.time() - t:.0f} s")
    its = s.iterations
    np.savez_compressed(
        os.path.join(HERE, "cfg2_trajectory.npz"),
        cost=np.array([i.cost for i in its]), successful=np.array([i.step_is_successful for i in its]),
        valid=np.array([i.step_is_valid for i in its]), radius=np.array([i.trust_region_radius for i in its]),
        step_norm=np.array([i.step_norm for i in its]), initial_cost=s.initial_cost, final_cost=s.final_cost,
        termination=s.termination, q=s.state.q, t=s.state.t, pt_index=np.arange(0, prob.num_points, 97),
        X=s.state.X[::97],
    )


def cfg3():
    prob = make_config("cfg3")
    C, P, N = prob.num_images, prob.num_points, prob.num_obs
    radius = 1e4
    t = time.time()
    st = State.of(prob)
    lin = evaluate(prob, st, jacobians=True)
    print(f"cfg3: linearised in {time.time() - t:.0f} s, cost {lin.cost!r}")
    freec = free_columns(prob)
    freep = np.repeat((prob.pt_const == 0)[:, None], 3, axis=1)
    ne_u = lm.NormalEquations(prob, lin, freec.astype(float), freep.astype(float))
    out = dict(
        obs_index=np.arange(0, N, 997), r_reproj=lin.r_reproj[::997], cost=lin.cost, fixed_cost=lin.fixed_cost,
        Hcc=ne_u.Hcc, gc=ne_u.gc, pt_index=np.arange(0, P, 1009), Hpp=ne_u.Hpp[::1009], gp=ne_u.gp[::1009],
    )
    dpos = np.full(N, -1, dtype=np.int64)
    dpos[prob.depth_obs] = np.arange(prob.depth_obs.shape[0])
    out["r_depth"] = lin.r_depth[dpos[::997]]
    n2c, n2p = ne_u.column_sq_norms()  # all blocks of cfg3 have free columns except the gauge
    ne0 = lm.NormalEquations(prob, lin, np.ones((C, 6)), np.ones((P, 3)))
    n2c, n2p = ne0.column_sq_norms()
    del ne0, ne_u
    c_img = np.where(freec, 1.0 / (1.0 + np.sqrt(n2c)), 0.0)
    c_pt = np.where(freep, 1.0 / (1.0 + np.sqrt(n2p)), 0.0)
    ne = lm.NormalEquations(prob, lin, c_img, c_pt)
    dc, dp = ne.column_sq_norms()
    Dc2, Dp2 = np.clip(dc, 1e-6, 1e32) / radius, np.clip(dp, 1e-6, 1e32) / radius
    S, rhs, _ = ne.schur_system(Dc2, Dp2)
    print(f"cfg3: reduced system at {time.time() - t:.0f} s")
    # unscaled variables, as the library reports them: S_u = diag(1/c) S diag(1/c) on free coordinates
    cs = c_img.reshape(-1)
    free = cs > 0
    inv = np.where(free, 1.0 / np.where(free, cs, 1.0), 0.0)
    rows = (6 * S_SAMPLE_IMAGES[:, None] + np.arange(6)[None, :]).reshape(-1)
    Ssub = S.tocsr()[rows][:, rows].toarray() * np.outer(inv[rows], inv[rows])
    out.update(s_images=S_SAMPLE_IMAGES, S_sample=Ssub, rhs=(rhs * inv).reshape(C, 6), radius=radius)
    yc, yp = ne.solve(Dc2, Dp2)
    out.update(y_c=yc * c_img, y_p=(yp * c_pt)[::1009])
    cand = lm.plus(prob, st, -yc * c_img, -yp * c_pt)
    cand_cost = evaluate(prob, cand, jacobians=False).cost
    oi, op = prob.obs_img, prob.obs_pt
    m = np.einsum("nki,ni->nk", ne.Jc, yc[oi]) + np.einsum("nki,ni->nk", ne.Jp, yp[op])
    mcc = float(np.sum(m * (ne.r - 0.5 * m)))
    rho = (lin.cost - cand_cost) / mcc
    out.update(candidate_cost=cand_cost, model_cost_change=mcc, relative_decrease=rho,
               cand_t=cand.t, cand_q=cand.q, cand_X=cand.X[::1009])
    print(f"cfg3: candidate cost {cand_cost!r}, rho {rho:.4f}, {time.time() - t:.0f} s")
    np.savez_compressed(os.path.join(HERE, "cfg3_first_iteration.npz"), **out)


if __name__ == "__main__":
    which = sys.argv[1:] or ["cfg2", "cfg3"]
    if "cfg2" in which:
        cfg2()
    if "cfg3" in which:
        cfg3()
