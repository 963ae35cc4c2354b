"""Every entry point of include/spt.h and every kernel variant once on tiny workloads, in a FRESH process (tools/sanitize_smoke.py —
also the workload for compute-sanitizer where that tool is open): all five configs, tiny pool (many iterations + the drain), unsorted
and counting kernels, box radius 1.5, normal AOV, general and lean shade kernels, padding-only iterations, ray batches incl. the
no-allocation one, a one-rank communicator created BEFORE torch is imported (the binding must load torch's NCCL first), film save/load,
device-built BLAS."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_every_entry_point_once_in_a_fresh_process(built):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "sanitize_smoke.py")], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "sanitize smoke done" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
