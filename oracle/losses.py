"""Ceres 2.1 loss functions and corrector (upstream ``internal/ceres/loss_function.cc``,
``corrector.cc``; not in /root/reference).  TEST INFRASTRUCTURE ONLY.

Reached from the reference through ``LossFunctionType`` at
/root/reference/mpsfm/sfm/mapper/bundle_adjustment.py:44-48 and :100; the fork wraps each loss in a
``ScaledLoss(loss, magnitude)`` (:99 ``loss_function_magnitude``; :161/:169 per-observation magnitudes).
"""

from __future__ import annotations

import numpy as np

TRIVIAL, SOFT_L1, CAUCHY, HUBER = 0, 1, 2, 3
_DBL_MIN = np.finfo(np.float64).tiny


def rho(loss_type: int, a, s):
    """(rho, rho', rho'') of ``loss_type`` with scale ``a`` at squared norm ``s`` (arrays)."""
    s = np.asarray(s, dtype=np.float64)
    a = np.broadcast_to(np.asarray(a, dtype=np.float64), s.shape)
    if loss_type == TRIVIAL:
        return s.copy(), np.ones_like(s), np.zeros_like(s)
    b = a * a
    c = 1.0 / b
    if loss_type == SOFT_L1:
        sm = 1.0 + s * c
        tmp = np.sqrt(sm)
        r1 = np.maximum(_DBL_MIN, 1.0 / tmp)
        return 2.0 * b * (tmp - 1.0), r1, -(c * r1) / (2.0 * sm)
    if loss_type == CAUCHY:
        sm = 1.0 + s * c
        inv = 1.0 / sm
        return b * np.log(sm), np.maximum(_DBL_MIN, inv), -c * (inv * inv)
    if loss_type == HUBER:
        outl = s > b
        r = np.sqrt(np.where(outl, s, 1.0))
        r1o = np.maximum(_DBL_MIN, a / r)
        return (
            np.where(outl, 2.0 * a * r - b, s),
            np.where(outl, r1o, 1.0),
            np.where(outl, -r1o / (2.0 * np.where(outl, s, 1.0)), 0.0),
        )
    raise ValueError(f"unknown loss {loss_type}")


def scaled_rho(loss_type: int, a, weight, s):
    """``ceres::ScaledLoss``: every component multiplied by ``weight``."""
    r0, r1, r2 = rho(loss_type, a, s)
    return weight * r0, weight * r1, weight * r2


def corrector(s, r1, r2):
    """``ceres::internal::Corrector``: returns (residual_scaling, alpha_sq_norm).

    residual <- residual_scaling * residual;
    J <- sqrt(rho') (J - alpha_sq_norm * r r^T J).
    For the losses above rho'' <= 0 always, which is Ceres' early-out branch (alpha = 0)."""
    sqrt_r1 = np.sqrt(r1)
    simple = (s == 0.0) | (r2 <= 0.0)
    safe_r1 = np.where(simple, 1.0, r1)
    D = 1.0 + 2.0 * s * r2 / safe_r1
    alpha = 1.0 - np.sqrt(np.maximum(D, 0.0))
    safe_s = np.where(simple, 1.0, s)
    scaling = np.where(simple, sqrt_r1, sqrt_r1 / np.where(simple, 1.0, 1.0 - alpha))
    alpha_sq = np.where(simple, 0.0, alpha / safe_s)
    return scaling, alpha_sq, sqrt_r1
