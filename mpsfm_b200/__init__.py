"""mpsfm_b200 -- B200-native bundle adjustment for MP-SfM.

Drop-in for the reference's pyceres-backed ``Optimizer``
(/root/reference/mpsfm/sfm/mapper/bundle_adjustment.py:18).  Host code is Python;
all arithmetic of the solve runs in hand-written sm_100a CUDA kernels reached through
the C-ABI declared in ``include/mpba.h`` (ctypes).  There is no CPU fallback: importing
the package works anywhere, but every compute entry point raises if ``libmpba.so`` or a
CUDA device is missing.
"""

from .problem import BundleProblem, CameraModel, LossType  # noqa: F401

__all__ = ["BundleProblem", "CameraModel", "LossType"]
