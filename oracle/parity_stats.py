"""Converged-image parity statistics: the CUDA core against the reference-faithful oracle mode — TEST INFRASTRUCTURE ONLY
(imported by tests/ and by the checker leg of bench.py that fills the "parity" key; never by the product package).

Bar (SURVEY.md §8c / BASELINE.md §4): relMSE(GPU, A) <= 2 x relMSE(A, B) and |mean-luminance ratio - 1| <= 0.5 %, where
A and B are two independent renders of the reference's algorithm at S spp and relMSE = mean((I - R)^2 / (R^2 + 1e-2)).

  A = reference-faithful in every respect: the reference's own BVH (Q1/Q2), one xoshiro stream and one stateful pixel sampler
      per thread, contiguous row bands (oracle rng_mode 0, fix_q1 off).
  B = the same algorithm on the checker BVH (fix_q1: same hits, ~30x faster); it only measures the noise floor.

The reference's estimator divides by max(pdf, 1e-5) (pt.rs:172-185), so on the conductor / glass scenes single samples reach 1e5..1e9
and the PLAIN mean does not converge at any affordable spp: at 4096 spp the mean luminance of two reference renders of C2 differs by
12x.  The plain numbers are reported, and the bar is checked on a bounded statistic of the SAME samples: the per-pixel mean of
min(sample, CLAMP) per channel (E[min(X, c)] is a functional of the sample distribution like the mean is, and both sides compute it
per sample: the oracle inside its pixel loop, the CUDA core by rendering the S samples of the frame as S one-sample passes through
spp_first / spp_count, which are the very same paths as the one-call frame)."""
import time

import numpy as np

CLAMP = 32.0  # above every directly visible emitter of C1..C4 (18 in C2), far below the estimator's outliers
CONVERGED = {  # name -> (width, height, spp)
    "c1": (64, 64, 4096),
    "c1b": (32, 18, 4096),
    "c2": (32, 18, 4096),
    "c3": (32, 18, 4096),
    "c4": (32, 18, 4096),
    "c5": (32, 18, 2048),
}
REF_THREADS = 18  # fixed, so that the per-thread streams of the faithful mode (and with them the test) do not depend on the host's core count


def relmse(x, r):
    x, r = np.asarray(x, np.float64), np.asarray(r, np.float64)
    return float((((x - r) ** 2) / (r ** 2 + 1e-2)).mean())


def lum(img):
    img = np.asarray(img, np.float64)
    return float((0.299 * img[..., 0] + 0.587 * img[..., 1] + 0.114 * img[..., 2]).mean())


def gpu_sample_statistics(sc, w, h, spp, clamp=CLAMP, seed=3, **kw):
    """(plain mean, mean of min(sample, clamp)) of the S samples per pixel, through S one-sample passes of the same frame."""
    acc, accc = np.zeros((h, w, 3), np.float64), np.zeros((h, w, 3), np.float64)
    for k in range(spp):
        img, _ = sc.render(w, h, spp=spp, spp_range=(k, 1), seed=seed, **kw)
        acc += img
        accc += np.minimum(img, np.float32(clamp))
    return (acc / spp).astype(np.float32), (accc / spp).astype(np.float32)


def converged_parity(sc, fs, name, size=None, clamp=CLAMP, check_one_call=True):
    """Runs A, B and the CUDA core; returns the dict that goes into the bench line / DESIGN.md."""
    import orc
    w, h, spp = size or CONVERGED[name]
    t0 = time.time()
    a_img, a_st = orc.OracleScene(fs, fix_q1=False).render(w, h, spp=spp, rng_mode=0, seed=1, n_threads=min(REF_THREADS, h), sample_clamp=clamp)
    t_a = time.time() - t0
    b_img, b_st = orc.OracleScene(fs, fix_q1=True).render(w, h, spp=spp, rng_mode=0, seed=2, n_threads=min(REF_THREADS, h), sample_clamp=clamp)
    t1 = time.time()
    g_img, g_cl = gpu_sample_statistics(sc, w, h, spp, clamp)
    t_g = time.time() - t1
    out = {"config": name, "width": w, "height": h, "spp": spp, "clamp": clamp,
           "reference_mode": "A: own BVH (Q1/Q2), per-thread xoshiro + stateful sampler, %d row bands; B: same on the checker BVH" % min(REF_THREADS, h),
           "relmse_gpu_vs_ref": relmse(g_cl, a_st["clamped_mean"]), "relmse_ref_vs_ref": relmse(b_st["clamped_mean"], a_st["clamped_mean"]),
           "lum_ratio": lum(g_cl) / lum(a_st["clamped_mean"]), "lum_ratio_ref_vs_ref": lum(b_st["clamped_mean"]) / lum(a_st["clamped_mean"]),
           "unclamped": {"relmse_gpu_vs_ref": relmse(g_img, np.nan_to_num(a_img)), "relmse_ref_vs_ref": relmse(np.nan_to_num(b_img), np.nan_to_num(a_img)),
                         "lum_ratio": lum(g_img) / max(lum(np.nan_to_num(a_img)), 1e-30), "lum_ratio_ref_vs_ref": lum(np.nan_to_num(b_img)) / max(lum(np.nan_to_num(a_img)), 1e-30),
                         "max_pixel_ref": float(np.nan_to_num(a_img).max()), "max_pixel_gpu": float(g_img.max())},
           "clamped_sample_share": float((g_img.astype(np.float64).mean() - g_cl.astype(np.float64).mean()) / max(g_img.astype(np.float64).mean(), 1e-30)),
           "seconds": {"ref_a": t_a, "gpu_passes": t_g}}
    if check_one_call:  # the S one-sample passes ARE the frame: their mean equals the one-call render up to the float sum
        one, _ = sc.render(w, h, spp=spp, seed=3)
        out["passes_vs_one_call_max_rel"] = float((np.abs(one - g_img) / (np.abs(one) + 1e-3)).max())
    return out
