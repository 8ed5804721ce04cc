"""Times the assembly statistics kernels (mpba_depth_prior_terms, mpba_segmented_median) on resident inputs next
to the reference's numpy on the host, at cfg3's size (5 M depth-prior samples, 1 000 images).

    python scripts/bench_stats.py > profiles/r01_stats_cfg3.json
"""

import ctypes as C
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mpsfm_b200 import engine  # noqa: E402


def gpu_ms(fn, reps=10):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    n, num_images = 5_000_000, 1000
    rng = np.random.default_rng(0)
    img_of = np.sort(rng.integers(0, num_images, n)).astype(np.int32)
    seg_ptr = np.searchsorted(img_of, np.arange(num_images + 1)).astype(np.int32)
    row3 = np.concatenate([rng.standard_normal((num_images, 3)) * 0.1 + [0, 0, 1], rng.standard_normal((num_images, 1))], 1)
    xyz = rng.standard_normal((n, 3)) * 2 + [0, 0, 30]
    z = np.einsum("nk,nk->n", row3[img_of, :3], xyz) + row3[img_of, 3]
    d = z * np.exp(rng.standard_normal(n) * 0.05)
    v = (0.05 * d) ** 2
    lib = engine.load_library()
    dev = torch.device("cuda", 0)
    t = {k: torch.from_numpy(a).to(dev) for k, a in dict(img_of=img_of, row3=row3, xyz=xyz, d=d, v=v, seg=seg_ptr).items()}
    out = {k: torch.empty(n, dtype=torch.float64, device=dev) for k in ("w", "a", "logratio", "whitened")}
    mask = torch.empty(n, dtype=torch.uint8, device=dev)
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def terms():
        rc = lib.mpba_depth_prior_terms(None, n, t["img_of"].data_ptr(), num_images, t["row3"].data_ptr(), t["xyz"].data_ptr(),
                                        t["d"].data_ptr(), t["v"].data_ptr(), 1.5, 0, 2.0, None, mask.data_ptr(),
                                        out["w"].data_ptr(), out["a"].data_ptr(), out["logratio"].data_ptr(),
                                        out["whitened"].data_ptr(), stream)
        assert rc == 0

    nbytes = C.c_size_t(0)
    lib.mpba_segmented_median_workspace_bytes(n, num_images, C.byref(nbytes))
    ws = torch.empty(nbytes.value + 256, dtype=torch.uint8, device=dev)
    ws_ptr = (ws.data_ptr() + 255) & ~255
    med = torch.empty(num_images, dtype=torch.float64, device=dev)
    mad = torch.empty(num_images, dtype=torch.float64, device=dev)
    one = torch.tensor([0, n], dtype=torch.int32, device=dev)

    def per_image_median():
        assert lib.mpba_segmented_median(None, n, out["logratio"].data_ptr(), num_images, t["seg"].data_ptr(), ws_ptr,
                                         nbytes.value, med.data_ptr(), None, stream) == 0

    def global_mad():
        assert lib.mpba_segmented_median(None, n, out["whitened"].data_ptr(), 1, one.data_ptr(), ws_ptr, nbytes.value,
                                         med.data_ptr(), mad.data_ptr(), stream) == 0

    res = {"n": n, "num_images": num_images, "gpu": torch.cuda.get_device_name(0)}
    ms = gpu_ms(terms)
    # 4 + 24 + 8 + 8 read, 1 + 4 * 8 written per sample
    res["depth_prior_terms"] = {"ms": ms, "GB_per_s": n * (44 + 33) / ms / 1e6}
    ms = gpu_ms(per_image_median)
    res["median_per_image"] = {"ms": ms, "GB_per_s": n * 8 * 8 / ms / 1e6, "bytes": "8 passes x 8 B"}
    ms = gpu_ms(global_mad)
    res["global_mad"] = {"ms": ms, "GB_per_s": n * 8 * 18 / ms / 1e6, "bytes": "(8 + 2 + 8) x 8 B"}

    # the reference's numpy on the host (bundle_adjustment.py:136-161, 239-241, 311-330)
    t0 = time.perf_counter()
    div = d / z
    m = (d > 0) & (div < 1.5) & (div > 1 / 1.5)
    w = d**2 * (1 / v.clip(1e-6, None))
    a = 2.0 * v**0.5 / d
    lr = np.log((z / d).clip(1e-6, None))
    wh = (np.log(d) - np.log(z)) / np.clip(v**0.5 / d, 1e-6, None)
    t1 = time.perf_counter()
    ref_med = np.array([np.median(lr[a_:b_]) for a_, b_ in zip(seg_ptr[:-1], seg_ptr[1:])])
    t2 = time.perf_counter()
    mu = np.median(wh)
    ref_mad = np.median(np.abs(wh - mu))
    t3 = time.perf_counter()
    res["numpy_host"] = {"terms_ms": (t1 - t0) * 1e3, "median_per_image_ms": (t2 - t1) * 1e3, "global_mad_ms": (t3 - t2) * 1e3}
    per_image_median()
    got_med = med.cpu().numpy().copy()
    global_mad()
    res["check"] = {"median_max_abs_diff": float(np.max(np.abs(got_med - ref_med))),
                    "mad_rel_diff": float(abs(mad.cpu().numpy()[0] - ref_mad) / ref_mad),
                    "mask_mismatch": int(np.sum(mask.cpu().numpy().astype(bool) != m))}
    print(json.dumps(res))


if __name__ == "__main__":
    main()
