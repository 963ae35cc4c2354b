"""spt_trace_batch_noalloc — the ray batch of seam B2 with its device scratch in a fixed file-scope array (no cudaMalloc on the call
path, no counterpart in the reference) — against the existing spt_trace_batch through the
C ABI, on two calls in a row.  The second call is SMALLER than the first and uses other rays: the shared scratch keeps the first call's
contents, and none of them may leak into the second result.  Bit-exact, like every other ray-batch comparison of two CUDA paths here.
Collected last on purpose (file name), and self-contained: it builds its own context and scene, so it passes alone in a fresh process
and after the other tests."""
import numpy as np
import pytest

import rays as R

pytestmark = pytest.mark.gpu


def test_noalloc_batch_equals_trace_batch_on_two_calls_in_a_row(built):
    from spt import abi, core, scenes
    ctx = core.Context(0)
    fs = scenes.load("c2")
    sc = core.Scene(ctx, fs)
    first = np.concatenate([R.primary_rays(fs, 192, 108), R.incoherent_rays(fs, abi.NOALLOC_MAX_RAYS - 192 * 108, seed=5)])
    assert len(first) == abi.NOALLOC_MAX_RAYS                     # exactly the limit
    second = R.incoherent_rays(fs, 12345, seed=6, finite_tmax=True)  # ragged size, other rays, finite t_max
    for rays in (first, second):
        want = sc.trace(rays)
        got = sc.trace_noalloc(rays)
        assert np.array_equal(got.view(np.uint8), want.view(np.uint8))
        assert 0.05 < (got["inst_id"] >= 0).mean() < 1.0
    again = sc.trace_noalloc(first)                               # and back to the big batch: same answer as before
    assert np.array_equal(again.view(np.uint8), sc.trace(first).view(np.uint8))
    assert len(sc.trace_noalloc(np.zeros(0, R.RAY_DT))) == 0      # empty batch
    with pytest.raises(abi.SptError, match="SPT_NOALLOC_MAX_RAYS"):
        sc.trace_noalloc(np.zeros(abi.NOALLOC_MAX_RAYS + 1, R.RAY_DT))
    sc.close()
    ctx.close()
