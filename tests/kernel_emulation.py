"""numpy emulation of what the CUDA kernels do with the per-observation quantities (N, h, p_rot, R).

TEST INFRASTRUCTURE: the per-observation arithmetic comes from the product's ``ba_math.cuh`` compiled
with g++ (``tests/hostcheck``); the reductions below follow ``ba_kernels.cu`` step by step so that the
algebraic restructuring used on the GPU (W = G^T N R, M = Ga^T (Qa Hpp^-1 Qb^T) Gb, the closed form
of the model cost change) is checked against the oracle's explicit Jacobians on the CPU box."""

import ctypes as C

import numpy as np


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def obs_quantities(hc, problem, state_q, state_t, state_X, want_jac=True):
    oi, op = problem.obs_img, problem.obs_pt
    n = oi.shape[0]
    pose = np.zeros((n, 8))
    pose[:, :4], pose[:, 4:7] = state_q[oi], state_t[oi]
    X = np.ascontiguousarray(state_X[op])
    cam = problem.img_camera[oi]
    model = np.ascontiguousarray(problem.cam_model[cam].astype(np.int32))
    prm = np.ascontiguousarray(problem.cam_params[cam])
    xy = np.ascontiguousarray(problem.obs_xy)
    dwa = np.zeros((n, 3))
    dwa[:, 0], dwa[:, 2] = 1.0, 1.0
    dwa[problem.depth_obs, 0] = problem.depth_d
    dwa[problem.depth_obs, 1] = problem.depth_w
    dwa[problem.depth_obs, 2] = problem.depth_a
    bs = np.ascontiguousarray(problem.shift_scale[oi])
    out = np.zeros((n, 48))
    plane = None
    if getattr(problem, "num_plane", 0):
        plane = np.zeros((n, 4))
        plane[problem.plane_obs, :3], plane[problem.plane_obs, 3] = problem.plane_m, problem.plane_c
    hc.hc_obs_plane(C.c_int(n), _p(pose), _p(X), _p(model), _p(prm), _p(xy), _p(dwa), _p(bs),
                    C.c_int(int(problem.reproj_loss)), C.c_int(int(problem.depth_loss)),
                    C.c_double(problem.reproj_loss_scale), C.c_double(problem.reproj_loss_weight),
                    C.c_int(1 if want_jac else 0), _p(plane) if plane is not None else None,
                    C.c_double(float(getattr(problem, "plane_loss_scale", 1.0))), _p(out))
    iu = np.triu_indices(6)
    Hcc = np.zeros((n, 6, 6))
    Hcc[:, iu[0], iu[1]] = out[:, 16:37]
    Hcc[:, iu[1], iu[0]] = out[:, 16:37]
    N6 = out[:, 0:6]
    Nm = np.empty((n, 3, 3))
    Nm[:, 0, 0], Nm[:, 0, 1], Nm[:, 0, 2] = N6[:, 0], N6[:, 1], N6[:, 2]
    Nm[:, 1, 0], Nm[:, 1, 1], Nm[:, 1, 2] = N6[:, 1], N6[:, 3], N6[:, 4]
    Nm[:, 2, 0], Nm[:, 2, 1], Nm[:, 2, 2] = N6[:, 2], N6[:, 4], N6[:, 5]
    return {"N": Nm, "h": out[:, 6:9], "prot": out[:, 9:12], "rho": out[:, 12], "r": out[:, 13:16],
            "Hcc": Hcc, "gc": out[:, 37:43], "rn": out[:, 43]}


def g_matrix(p):
    n = p.shape[0]
    G = np.zeros((n, 3, 6))
    G[:, 0, 1], G[:, 0, 2] = 2 * p[:, 2], -2 * p[:, 1]
    G[:, 1, 0], G[:, 1, 2] = -2 * p[:, 2], 2 * p[:, 0]
    G[:, 2, 0], G[:, 2, 1] = 2 * p[:, 1], -2 * p[:, 0]
    G[:, 0, 3] = G[:, 1, 4] = G[:, 2, 5] = 1.0
    return G
