"""Ceres 2.1 trust-region Levenberg-Marquardt with an exact Schur solve.  TEST INFRASTRUCTURE ONLY.

Restates what ``pyceres.solve`` does for the reference
(/root/reference/mpsfm/sfm/mapper/bundle_adjustment.py:285-293: ``SPARSE_SCHUR``, otherwise Ceres
defaults) following upstream ceres-solver 2.1.0 ``internal/ceres/trust_region_minimizer.cc``,
``levenberg_marquardt_strategy.cc``, ``trust_region_step_evaluator.cc`` and
``schur_complement_solver.cc`` (not in /root/reference; docker/install_ceres_solver.sh:12).
PARITY UNPINNED against the real library -- see ``oracle/__init__.py``.

Order of events per iteration (Ceres ``TrustRegionMinimizer::Minimize``):
  ComputeTrustRegionStep -> (invalid? shrink, continue) -> candidate cost ->
  ParameterToleranceReached -> FunctionToleranceReached -> accept/reject -> radius update ->
  FinalizeIteration (max-iterations, gradient tolerance, min radius).
"""

from __future__ import annotations

import dataclasses

import numpy as np
import scipy.linalg
import scipy.sparse as sp
import scipy.sparse.linalg

from . import geometry
from .residuals import State, evaluate, free_columns


@dataclasses.dataclass
class Options:
    max_num_iterations: int = 50
    initial_trust_region_radius: float = 1e4
    max_trust_region_radius: float = 1e16
    min_trust_region_radius: float = 1e-32
    min_relative_decrease: float = 1e-3
    min_lm_diagonal: float = 1e-6
    max_lm_diagonal: float = 1e32
    function_tolerance: float = 1e-6
    gradient_tolerance: float = 1e-10
    parameter_tolerance: float = 1e-8
    jacobi_scaling: bool = True
    max_num_consecutive_invalid_steps: int = 5


@dataclasses.dataclass
class IterationSummary:
    iteration: int
    cost: float  # includes fixed_cost, like ceres::IterationSummary::cost
    cost_change: float = 0.0
    gradient_max_norm: float = 0.0
    step_norm: float = 0.0
    relative_decrease: float = 0.0
    trust_region_radius: float = 0.0
    step_is_valid: bool = False
    step_is_successful: bool = False


@dataclasses.dataclass
class Summary:
    iterations: list
    initial_cost: float
    final_cost: float
    fixed_cost: float
    termination: str
    message: str
    num_successful_steps: int
    num_unsuccessful_steps: int
    state: State


def segment_sum(idx, vals, n):
    """sum rows of vals[N,...] by key idx[N] -> [n,...] (bincount per trailing element)."""
    flat = vals.reshape(vals.shape[0], -1)
    out = np.empty((n, flat.shape[1]))
    for k in range(flat.shape[1]):
        out[:, k] = np.bincount(idx, weights=flat[:, k], minlength=n)
    return out.reshape((n,) + vals.shape[1:])


class NormalEquations:
    """Block form of J'^T J' for the scaled, robustified Jacobian J' = J diag(c)."""

    def __init__(self, problem, lin, c_img, c_pt):
        oi, op = problem.obs_img, problem.obs_pt
        C, P = problem.img_q.shape[0], problem.pt_xyz.shape[0]
        self.problem, self.C, self.P = problem, C, P
        self.Jc = lin.Jc * c_img[oi][:, None, :]
        self.Jp = lin.Jp * c_pt[op][:, None, :]
        self.r = lin.r
        self.Hcc = segment_sum(oi, np.einsum("nki,nkj->nij", self.Jc, self.Jc), C)
        self.Hpp = segment_sum(op, np.einsum("nki,nkj->nij", self.Jp, self.Jp), P)
        self.gc = segment_sum(oi, np.einsum("nki,nk->ni", self.Jc, self.r), C)
        self.gp = segment_sum(op, np.einsum("nki,nk->ni", self.Jp, self.r), P)
        self.W = np.einsum("nki,nkj->nij", self.Jc, self.Jp)  # [N,6,3]

    def column_sq_norms(self):
        return np.einsum("cii->ci", self.Hcc).copy(), np.einsum("pii->pi", self.Hpp).copy()

    def schur_system(self, Dc2, Dp2):
        """Reduced camera system S, rhs and the pieces back-substitution needs."""
        problem, C, P = self.problem, self.C, self.P
        oi, op = problem.obs_img, problem.obs_pt
        Hpp = self.Hpp + np.einsum("pi,ij->pij", Dp2, np.eye(3))
        Hinv = np.linalg.inv(Hpp)
        N = oi.shape[0]
        rows = (6 * oi[:, None, None] + np.arange(6)[None, :, None] + np.zeros((1, 1, 3), np.int64)).reshape(-1)
        cols = (3 * op[:, None, None] + np.arange(3)[None, None, :] + np.zeros((1, 6, 1), np.int64)).reshape(-1)
        Wm = sp.coo_matrix((self.W.reshape(-1), (rows, cols)), shape=(6 * C, 3 * P)).tocsr()
        Him = sp.bsr_matrix((Hinv, np.arange(P), np.arange(P + 1)), shape=(3 * P, 3 * P)).tocsr()
        WH = Wm @ Him
        Hcc = self.Hcc + np.einsum("ci,ij->cij", Dc2, np.eye(6))
        Hcm = sp.bsr_matrix((Hcc, np.arange(C), np.arange(C + 1)), shape=(6 * C, 6 * C)).tocsr()
        S = Hcm - WH @ Wm.T
        rhs = self.gc.reshape(-1) - WH @ self.gp.reshape(-1)
        del N
        return S, rhs, Hinv

    def solve(self, Dc2, Dp2):
        """Solve (J'^T J' + D^2) y = J'^T r exactly by Schur elimination of the points."""
        problem, C = self.problem, self.C
        oi, op = problem.obs_img, problem.obs_pt
        S, rhs, Hinv = self.schur_system(Dc2, Dp2)
        free = np.flatnonzero(free_columns(problem).reshape(-1))
        yc = np.zeros(6 * C)
        if free.size:
            Sff = S[free][:, free]
            if free.size <= 6000:
                cf = scipy.linalg.cho_factor(Sff.toarray(), lower=True)
                yc[free] = scipy.linalg.cho_solve(cf, rhs[free])
            else:
                yc[free] = scipy.sparse.linalg.splu(Sff.tocsc()).solve(rhs[free])
        yc = yc.reshape(C, 6)
        wty = segment_sum(op, np.einsum("nij,ni->nj", self.W, yc[oi]), self.P)
        yp = np.einsum("pij,pj->pi", Hinv, self.gp - wty)
        return yc, yp

    def dense(self, Dc2, Dp2):
        """Full dense (J'^T J' + D^2) and J'^T r over [images | points] -- for cross-checks on tiny scenes."""
        problem, C, P = self.problem, self.C, self.P
        oi, op = problem.obs_img, problem.obs_pt
        n = 6 * C + 3 * P
        H = np.zeros((n, n))
        for c in range(C):
            H[6 * c : 6 * c + 6, 6 * c : 6 * c + 6] = self.Hcc[c] + np.diag(Dc2[c])
        for p in range(P):
            o = 6 * C + 3 * p
            H[o : o + 3, o : o + 3] = self.Hpp[p] + np.diag(Dp2[p])
        for k in range(oi.shape[0]):
            a, b = 6 * oi[k], 6 * C + 3 * op[k]
            H[a : a + 6, b : b + 3] += self.W[k]
            H[b : b + 3, a : a + 6] += self.W[k].T
        g = np.concatenate([self.gc.reshape(-1), self.gp.reshape(-1)])
        return H, g


def _block_masks(problem):
    """Which ambient parameter blocks are in the reduced program (non-constant and referenced)."""
    C, P = problem.img_q.shape[0], problem.pt_xyz.shape[0]
    m = np.asarray(problem.img_free_mask, dtype=np.uint8)
    img_used = np.bincount(problem.obs_img, minlength=C) > 0
    q_act = ((m & 0b000111) != 0) & img_used
    t_act = ((m & 0b111000) != 0) & img_used
    p_act = (np.asarray(problem.pt_const) == 0) & (np.bincount(problem.obs_pt, minlength=P) > 0)
    return q_act, t_act, p_act


def _ambient_sq_norm(problem, st: State):
    q_act, t_act, p_act = _block_masks(problem)
    return float(np.sum(st.q[q_act] ** 2) + np.sum(st.t[t_act] ** 2) + np.sum(st.X[p_act] ** 2))


def _ambient_sq_dist(problem, a: State, b: State):
    q_act, t_act, p_act = _block_masks(problem)
    return float(
        np.sum((a.q[q_act] - b.q[q_act]) ** 2) + np.sum((a.t[t_act] - b.t[t_act]) ** 2) + np.sum((a.X[p_act] - b.X[p_act]) ** 2)
    )


def plus(problem, st: State, d_img, d_pt) -> State:
    """x (+) delta with the reference's manifolds (bundle_adjustment.py:114-122)."""
    return State(geometry.quat_plus(st.q, d_img[:, :3]), st.t + d_img[:, 3:], st.X + d_pt)


def solve(problem, options: Options | None = None, state: State | None = None, verbose: bool = False) -> Summary:
    opt = options or Options()
    x = (state or State.of(problem)).copy()
    oi, op = problem.obs_img, problem.obs_pt
    C, P = problem.img_q.shape[0], problem.pt_xyz.shape[0]
    freec = free_columns(problem)
    freep = (np.asarray(problem.pt_const) == 0)[:, None] & np.ones((1, 3), bool)

    # ---- IterationZero
    x_norm = np.sqrt(_ambient_sq_norm(problem, x))
    lin = evaluate(problem, x, jacobians=True)
    x_cost, fixed_cost = lin.cost, lin.fixed_cost
    c_img, c_pt = np.ones((C, 6)), np.ones((P, 3))
    ne = NormalEquations(problem, lin, c_img, c_pt)  # unscaled: gradient and iteration-0 column norms
    gmax = max(np.abs(ne.gc).max(initial=0.0), np.abs(ne.gp).max(initial=0.0))
    if opt.jacobi_scaling:
        n2c, n2p = ne.column_sq_norms()
        c_img = np.where(freec, 1.0 / (1.0 + np.sqrt(n2c)), 0.0)
        c_pt = np.where(freep, 1.0 / (1.0 + np.sqrt(n2p)), 0.0)
    else:
        c_img, c_pt = freec.astype(float), freep.astype(float)
    ne = NormalEquations(problem, lin, c_img, c_pt)

    radius, decrease_factor = opt.initial_trust_region_radius, 2.0
    its = [IterationSummary(0, x_cost + fixed_cost, gradient_max_norm=gmax, trust_region_radius=radius)]
    minimum_cost, best = x_cost, x.copy()
    n_ok = n_bad = 0
    invalid_run = 0
    term, msg = None, ""

    def finalize_check(it: IterationSummary):
        # FinalizeIterationAndCheckIfMinimizerCanContinue
        if it.iteration >= opt.max_num_iterations:
            return "NO_CONVERGENCE", "Maximum number of iterations reached."
        if it.gradient_max_norm <= opt.gradient_tolerance:
            return "CONVERGENCE", "Gradient tolerance reached."
        if it.trust_region_radius <= opt.min_trust_region_radius:
            return "CONVERGENCE", "Minimum trust region radius reached."
        return None

    chk = finalize_check(its[0])
    if chk:
        term, msg = chk
    reuse_diagonal = False
    diag_c = diag_p = None
    while term is None:
        it = IterationSummary(its[-1].iteration + 1, its[-1].cost, gradient_max_norm=its[-1].gradient_max_norm)
        # ---- LevenbergMarquardtStrategy::ComputeStep
        if not reuse_diagonal:
            diag_c, diag_p = ne.column_sq_norms()
            diag_c = np.clip(diag_c, opt.min_lm_diagonal, opt.max_lm_diagonal)
            diag_p = np.clip(diag_p, opt.min_lm_diagonal, opt.max_lm_diagonal)
        reuse_diagonal = True
        try:
            yc, yp = ne.solve(diag_c / radius, diag_p / radius)
            ok = bool(np.all(np.isfinite(yc)) and np.all(np.isfinite(yp)))
        except (np.linalg.LinAlgError, scipy.linalg.LinAlgError):
            ok = False
        model_cost_change = 0.0
        if ok:
            m = np.einsum("nki,ni->nk", ne.Jc, yc[oi]) + np.einsum("nki,ni->nk", ne.Jp, yp[op])  # = -J' step
            model_cost_change = float(np.sum(m * (ne.r - 0.5 * m)))
            ok = model_cost_change > 0.0
        it.step_is_valid = ok
        if not ok:
            # HandleInvalidStep
            invalid_run += 1
            radius /= decrease_factor
            decrease_factor *= 2.0
            it.trust_region_radius = radius
            n_bad += 1
            its.append(it)
            if invalid_run >= opt.max_num_consecutive_invalid_steps:
                term, msg = "FAILURE", "Number of consecutive invalid steps more than max_num_consecutive_invalid_steps."
                break
            chk = finalize_check(it)
            if chk:
                term, msg = chk
            continue
        invalid_run = 0
        d_img, d_pt = -yc * c_img, -yp * c_pt
        cand = plus(problem, x, d_img, d_pt)
        cand_cost = evaluate(problem, cand, jacobians=False).cost
        if not np.isfinite(cand_cost):
            cand_cost = np.finfo(np.float64).max
        # ---- ParameterToleranceReached
        it.step_norm = np.sqrt(_ambient_sq_dist(problem, x, cand))
        if it.step_norm <= opt.parameter_tolerance * (x_norm + opt.parameter_tolerance):
            term, msg = "CONVERGENCE", "Parameter tolerance reached."
            its.append(it)
            break
        # ---- FunctionToleranceReached
        it.cost_change = x_cost - cand_cost
        if abs(it.cost_change) <= opt.function_tolerance * x_cost:
            term, msg = "CONVERGENCE", "Function tolerance reached."
            its.append(it)
            break
        # ---- IsStepSuccessful (monotonic steps)
        it.relative_decrease = it.cost_change / model_cost_change
        if it.relative_decrease > opt.min_relative_decrease:
            x, x_cost = cand, cand_cost
            x_norm = np.sqrt(_ambient_sq_norm(problem, x))
            lin = evaluate(problem, x, jacobians=True)
            x_cost = lin.cost
            ne_u = NormalEquations(problem, lin, freec.astype(float), freep.astype(float))
            it.gradient_max_norm = max(np.abs(ne_u.gc).max(initial=0.0), np.abs(ne_u.gp).max(initial=0.0))
            ne = NormalEquations(problem, lin, c_img, c_pt)
            it.cost = x_cost + fixed_cost
            it.step_is_successful = True
            radius = min(opt.max_trust_region_radius, radius / max(1.0 / 3.0, 1.0 - (2.0 * it.relative_decrease - 1.0) ** 3))
            decrease_factor = 2.0
            reuse_diagonal = False
            n_ok += 1
            if x_cost < minimum_cost:
                minimum_cost, best = x_cost, x.copy()
        else:
            radius /= decrease_factor
            decrease_factor *= 2.0
            n_bad += 1
        it.trust_region_radius = radius
        its.append(it)
        if verbose:
            print(f"{it.iteration:3d} cost {it.cost:.9e} dcost {it.cost_change:.3e} |g| {it.gradient_max_norm:.3e} "
                  f"|step| {it.step_norm:.3e} rho {it.relative_decrease:.3e} radius {it.trust_region_radius:.3e}")
        chk = finalize_check(it)
        if chk:
            term, msg = chk
    return Summary(
        iterations=its,
        initial_cost=its[0].cost,
        final_cost=minimum_cost + fixed_cost,
        fixed_cost=fixed_cost,
        termination=term,
        message=msg,
        num_successful_steps=n_ok,
        num_unsuccessful_steps=n_bad,
        state=best,
    )
