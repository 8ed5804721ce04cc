"""Vectorised forward-mode dual numbers -- the oracle's stand-in for ``ceres::Jet``.

The reference differentiates its cost functors with Ceres autodiff (pycolmap cost functors are
``ceres::AutoDiffCostFunction``; call site /root/reference/mpsfm/sfm/mapper/bundle_adjustment.py:104).
A ``Jet`` carries a value array ``a[n]`` and a derivative array ``v[n,D]``.  Test infrastructure only.
"""

from __future__ import annotations

import numpy as np


class Jet:
    __slots__ = ("a", "v")
    __array_priority__ = 1000
    __array_ufunc__ = None  # make ndarray <op> Jet defer to Jet.__r<op>__

    def __init__(self, a, v):
        self.a = np.asarray(a, dtype=np.float64)
        self.v = np.asarray(v, dtype=np.float64)

    @staticmethod
    def variables(values):
        """values [n,D] -> list of D jets seeded with the identity."""
        values = np.asarray(values, dtype=np.float64)
        n, D = values.shape
        out = []
        for j in range(D):
            v = np.zeros((n, D))
            v[:, j] = 1.0
            out.append(Jet(values[:, j].copy(), v))
        return out

    # -- helpers
    def _lift(self, o):
        if isinstance(o, Jet):
            return o
        o = np.asarray(o, dtype=np.float64)
        return Jet(np.broadcast_to(o, self.a.shape), np.zeros_like(self.v))

    def __neg__(self):
        return Jet(-self.a, -self.v)

    def __add__(self, o):
        o = self._lift(o)
        return Jet(self.a + o.a, self.v + o.v)

    __radd__ = __add__

    def __sub__(self, o):
        o = self._lift(o)
        return Jet(self.a - o.a, self.v - o.v)

    def __rsub__(self, o):
        return self._lift(o) - self

    def __mul__(self, o):
        o = self._lift(o)
        return Jet(self.a * o.a, self.a[:, None] * o.v + o.a[:, None] * self.v)

    __rmul__ = __mul__

    def __truediv__(self, o):
        o = self._lift(o)
        inv = 1.0 / o.a
        q = self.a * inv
        return Jet(q, (self.v - q[:, None] * o.v) * inv[:, None])

    def __rtruediv__(self, o):
        return self._lift(o) / self


def _un(x, f, df):
    if isinstance(x, Jet):
        return Jet(f(x.a), df(x.a)[:, None] * x.v)
    return f(np.asarray(x, dtype=np.float64))


def sqrt(x):
    return _un(x, np.sqrt, lambda a: 0.5 / np.sqrt(a))


def atan(x):
    return _un(x, np.arctan, lambda a: 1.0 / (1.0 + a * a))


def tan(x):
    return _un(x, np.tan, lambda a: 1.0 + np.tan(a) ** 2)


def log(x):
    return _un(x, np.log, lambda a: 1.0 / a)


def value(x):
    return x.a if isinstance(x, Jet) else np.asarray(x, dtype=np.float64)


def where(mask, x, y):
    """Elementwise select; mask is a plain bool array (branches on the value part, like Ceres)."""
    if isinstance(x, Jet) or isinstance(y, Jet):
        ref = x if isinstance(x, Jet) else y
        x, y = ref._lift(x), ref._lift(y)
        return Jet(np.where(mask, x.a, y.a), np.where(mask[:, None], x.v, y.v))
    return np.where(mask, x, y)
