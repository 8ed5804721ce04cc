"""Configuration handling of the drop-in classes.

The reference merges a per-class ``default_conf`` dict with the user's OmegaConf node and freezes the
top level (``BaseClass.__init__``, /root/reference/mpsfm/baseclass.py:18-31 and
``freeze_top_level_cfg``, /root/reference/mpsfm/utils/tools.py:95-99): unknown top-level keys raise, nested
dicts stay open.  OmegaConf is used when it is importable (it is wherever the reference runs); otherwise
``Conf`` below reproduces exactly that behaviour for plain dicts, so the package has no hard dependency.
"""

from __future__ import annotations

from time import time


class Conf(dict):
    """Attribute-style dict; ``_frozen`` instances reject unknown keys (struct mode)."""

    _frozen = False

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:
            raise AttributeError(k) from e

    def __setattr__(self, k, v):
        if k == "_frozen":
            object.__setattr__(self, k, v)
            return
        if self._frozen and k not in self:
            raise AttributeError(f"Key '{k}' is not in struct")
        self[k] = v


def _to_conf(d) -> Conf:
    out = Conf()
    for k, v in dict(d).items():
        out[k] = _to_conf(v) if isinstance(v, dict) else v
    return out


def merge_conf(default_conf: dict, conf=None):
    """default_conf <- conf with the reference's struct semantics."""
    try:
        from omegaconf import DictConfig, OmegaConf
    except ImportError:
        OmegaConf = None
    if OmegaConf is not None:
        base = OmegaConf.create(default_conf)
        OmegaConf.set_struct(base, True)
        for k in base:
            if isinstance(base[k], DictConfig):
                OmegaConf.set_struct(base[k], False)
        if conf is None:
            conf = {}
        if isinstance(conf, dict):
            conf = OmegaConf.create(conf)
        return OmegaConf.merge(base, conf)
    base = _to_conf(default_conf)
    user = {} if conf is None else (conf if isinstance(conf, dict) else dict(conf))
    for k, v in user.items():
        if k not in base:
            raise AttributeError(f"Key '{k}' is not in struct (valid keys: {sorted(base)})")
        if isinstance(base[k], Conf) and isinstance(v, dict):
            base[k].update(_to_conf(v))
        else:
            base[k] = v
    base._frozen = True
    return base


class BaseClass:
    """Mirror of the reference's ``BaseClass`` (baseclass.py:8-51): conf merge, ``_init`` hook, ``log``."""

    default_conf = {"verbose": 0}

    def __init__(self, conf=None, *args, **kwargs):
        self.conf = merge_conf(self.default_conf, conf)
        self._assert_configs()
        self._propagate_conf()
        self._init(*args, **kwargs)
        self.tstart = None

    def _init(self, *args, **kwargs):
        pass

    def _assert_configs(self):
        pass

    def _propagate_conf(self):
        pass

    def log(self, *message, level=0, tstart=False, tend=False, **kwargs):
        if self.conf.verbose >= level:
            if tstart:
                self.tstart = time()
            elif tend:
                assert len(message) == 0, "Message should be None when tend is True"
                assert self.tstart is not None, "tstart should be set before tend"
                message = [f"{time() - self.tstart:.3f} s"]
            print(*message, end=" " if tstart else "\n", **kwargs)
