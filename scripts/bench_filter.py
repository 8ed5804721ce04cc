"""Times the post-BA track filter (mpba_filter_tracks) on resident arrays at cfg3's size
(1 000 images / 1 M points / 5 M observations, SIMPLE_RADIAL).

    python scripts/bench_filter.py > profiles/r01_filter_cfg3.json

The call time includes everything the C-ABI call does (pose normalisation, staging the points, the filter kernel, the
read-back of the filtered count); algorithmic bytes = 21 B per observation (image index, keypoint, keep flag) + 45 B per
point (xyz record, track pointer, keep flag, error).
"""

import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mpsfm_b200 import engine  # noqa: E402
from mpsfm_b200.problem import CameraModel  # noqa: E402
from mpsfm_b200.synthetic import make_scene  # noqa: E402


def main():
    C_, P, L = 1000, 1_000_000, 5
    prob = make_scene(C_, P, L, model=CameraModel.SIMPLE_RADIAL, seed=0, outlier_frac=0.05, noise_px=0.7,
                      pose_rot_noise=0.0, pose_pos_noise=0.0, point_noise=0.002, depth_prior=False)
    ba = engine.BundleAdjuster(prob, device=0)
    ok, pk, pe, nf = ba.filter_tracks(4.0, 1.5)
    lay = ba.layout
    tk = torch.empty(lay.N, dtype=torch.uint8, device=ba.device)
    tp = torch.empty(lay.P, dtype=torch.uint8, device=ba.device)
    te = torch.empty(lay.P, dtype=torch.float64, device=ba.device)
    ws, wsn = ba._ws()
    import ctypes as C

    def call():
        rc = ba.lib.mpba_filter_tracks(ba.handle, C.byref(ba.prob), ws, wsn, 4.0, 1.5, None, tk.data_ptr(), tp.data_ptr(),
                                       te.data_ptr(), None)
        assert rc == 0

    for _ in range(3):
        call()
    torch.cuda.synchronize()
    reps = 20
    t0 = time.perf_counter()
    for _ in range(reps):
        call()
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) / reps * 1e3
    nbytes = 21 * lay.N + 45 * lay.P
    print(json.dumps({
        "what": "mpba_filter_tracks, resident arrays, whole C-ABI call (normalise poses + stage points + filter + count read-back)",
        "images": C_, "points": P, "observations": int(lay.N), "max_reproj_error_px": 4.0, "min_tri_angle_deg": 1.5,
        "observations_filtered": int(nf), "points_removed": int((~pk).sum()),
        "call_ms": ms, "observations_per_s": lay.N / ms * 1e3,
        "algorithmic_bytes": nbytes, "algorithmic_GBps_of_the_call": nbytes / ms / 1e6,
        "count_matches_flags": bool(nf == int((~ok).sum())),
    }))


if __name__ == "__main__":
    main()
