"""oracle/integration.py (the numpy restatement of the depth-integration solver) against golden vectors recorded from
the reference's own Integration class (tests/golden/make_golden_integration.py).  Runs anywhere (no GPU)."""

import os

import numpy as np
import pytest

from oracle import integration as oi

GOLD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "integration.npz"))
CASES = [str(c) for c in GOLD["cases"]]


def load_case(name):
    inp = {k.split("/")[-1]: GOLD[k] for k in GOLD.files if k.startswith(f"{name}/in/")}
    out = {k.split("/")[-1]: GOLD[k] for k in GOLD.files if k.startswith(f"{name}/out/")}
    over = {k.split("/")[-1]: GOLD[k].item() for k in GOLD.files if k.startswith(f"{name}/conf/")}
    conf = oi.IntegrationConf(**{k: v for k, v in over.items() if hasattr(oi.IntegrationConf, k)})
    return inp, out, conf


def prepared(inp, conf, depth_init=None, prior=None):
    return oi.prepare(inp["depth_prior"] if prior is None else prior, inp["uncertainty"], inp["valid"], inp["normals"],
                      inp["normals_uncertainty"], inp["depth_init"] if depth_init is None else depth_init, inp["kps"],
                      inp["depth3d"], inp["zvars3d"], inp["K"], conf)


@pytest.mark.parametrize("name", CASES)
def test_first_integration_matches_the_reference(name):
    inp, out, conf = load_case(name)
    trace = {}
    ok, depth, state = oi.integrate(prepared(inp, conf), conf, trace=trace)
    assert ok == bool(out["ok"])
    assert list(trace["cg_iterations"]) == list(out["cg_iterations"])  # same CG iteration counts in every IRLS step
    assert np.allclose(trace["energies"], out["energies"], rtol=1e-9)
    for za, zb in zip(trace["z_steps"], out["z_steps"]):
        assert np.max(np.abs(za - zb)) <= 1e-9 * np.max(np.abs(zb))
    assert np.max(np.abs(depth - out["depth"]) / out["depth"]) <= 1e-9
    assert np.allclose(state.wu, out["wu"], rtol=0, atol=1e-9) and np.allclose(state.wv, out["wv"], rtol=0, atol=1e-9)
    assert abs(state.energy_old - float(out["energy_old"])) <= 1e-9 * float(out["energy_old"])


@pytest.mark.parametrize("name", [c for c in CASES if f"{c}/out/second_ok" in GOLD.files])
def test_second_call_uses_the_cached_state_like_the_reference(name):
    inp, out, conf = load_case(name)
    state = oi.IntegrationState(True, float(out["energy_old"]), out["wu"], out["wv"])
    trace = {}
    pr = prepared(inp, conf, depth_init=out["second_depth_init"], prior=out["second_prior"])
    ok, depth, st2 = oi.integrate(pr, conf, state, trace=trace)
    assert ok == bool(out["second_ok"])
    assert list(trace["cg_iterations"]) == list(out["second_cg_iterations"])
    assert np.allclose(trace["energies"], out["second_energies"], rtol=1e-9)
    if ok:
        assert np.max(np.abs(depth - out["second_depth"]) / out["second_depth"]) <= 1e-9
    assert abs(st2.energy_old - float(out["second_energy_old"])) <= 1e-9 * float(out["second_energy_old"])


def test_hessian_stencil_and_propagated_variances():
    name = "step_36x48"
    inp, out, conf = load_case(name)
    H, W = inp["depth_prior"].shape
    import dataclasses

    for tag, ign in (("ign", True), ("full", False)):
        # calculate_hessian takes the sparse depths unfiltered (no scale filter there)
        pr = prepared(inp, dataclasses.replace(conf, scale_filter=False), depth_init=out["hess_depth"], prior=out["second_prior"])
        diag, e_r, e_d = oi.hessian(pr, conf, ignore_depths=ign)
        assert np.allclose(diag, out[f"hess_{tag}_diag"], rtol=1e-12)
        assert np.allclose(e_r, out[f"hess_{tag}_er"], rtol=1e-12, atol=0) and np.allclose(e_d, out[f"hess_{tag}_ed"], rtol=1e-12)
        var = oi.variances_at(H, W, diag, e_r, e_d, out["var_query"])
        # the reference's cholespy stub solved exactly in double, cast through a float32 torch tensor (CholeskySolverF)
        assert np.allclose(var, out[f"var_{tag}"], rtol=2e-6)


def test_stencil_operators_are_adjoint_and_the_matvec_is_the_normal_matrix():
    rng = np.random.default_rng(0)
    H, W = 9, 13
    n = H * W
    st = oi.Stencil(H, W, rng.normal(size=n), rng.normal(size=n))
    z, v = rng.normal(size=n), [rng.normal(size=n) for _ in range(4)]
    a = st.apply(z)
    assert np.isclose(sum(np.dot(ai, vi) for ai, vi in zip(a, v)), np.dot(z, st.transpose_sum(*v)))
    w = [rng.random(n) for _ in range(4)]
    prec = rng.random(n)
    diag, e_r, e_d = oi.normal_equations(H, W, st.nu.reshape(-1), st.nv.reshape(-1), *w, prec, 1.0, np.zeros(0, int),
                                         np.zeros(0), 1.0)
    lhs = oi.stencil_matvec(H, W, diag, e_r, e_d, z)
    rhs = st.transpose_sum(*[wi * ai for wi, ai in zip(w, a)]) + prec * z
    assert np.allclose(lhs, rhs, rtol=1e-12, atol=1e-12)
