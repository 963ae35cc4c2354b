"""Host logic of bench.py that decides what the CPU arm measures (no GPU): the bounded sample must keep the reference's row bands
balanced — PathTracer::render gives each of 2 x cores threads height / threads rows and the LAST thread the remainder
(renderer/util.rs:6-19), and its wall time is the slowest band's; round 1 sampled 160 x 90, which left 28 of 90 rows on one thread."""
import importlib.util
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("bench", os.path.join(ROOT, "bench.py"))
bench = importlib.util.module_from_spec(spec)
spec.loader.exec_module(bench)


@pytest.mark.parametrize("cores", [1, 4, 8, 16, 32, 64, 96])
@pytest.mark.parametrize("wh", [(1920, 1080), (512, 512), (1280, 720), (3840, 2160)])
def test_cpu_sample_keeps_the_row_bands_balanced(cores, wh):
    w, h = wh
    sw, sh, threads = bench.cpu_sample_size(w, h, cores)
    assert threads == 2 * cores and 1 <= sh <= h and sw >= 16
    t = min(threads, sh)
    per = sh // t
    last = sh - per * (t - 1)
    assert per >= 1 and last == per, (sw, sh, per, last)         # every band the same height: no thread carries a remainder
    if 4 * threads <= h and threads >= 8:
        assert per == 4                                             # four rows per band whenever the frame is tall enough
    assert abs(sw / sh - w / h) / (w / h) < 0.05                    # same aspect ratio as the full configuration
    assert "rows per band" in bench.band_note(sh, threads)


def test_relmse_and_luminance_helpers():
    import sys
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import parity_stats as PS
    a = np.full((4, 4, 3), 0.5, np.float32)
    assert PS.relmse(a, a) == 0.0 and abs(PS.lum(a) - 0.5) < 1e-7     # 0.299 + 0.587 + 0.114 = 1
    b = a.copy(); b[0, 0, 0] = 1.0
    assert abs(PS.relmse(b, a) - (0.25 / (0.25 + 1e-2)) / 48) < 1e-9
    assert set(PS.CONVERGED) >= {"c1", "c2", "c3", "c4", "c5"} and all(spp >= 2048 for _, _, spp in PS.CONVERGED.values())
