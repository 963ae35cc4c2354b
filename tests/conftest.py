import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "simple-path-tracer_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def built():
    import __graft_entry__ as g
    return g.build()


@pytest.fixture(scope="session")
def ctx(built):
    from spt import core
    c = core.Context(0)
    yield c
    c.close()


_scene_cache = {}


@pytest.fixture(scope="session")
def scene_pair(ctx):
    """(name) -> (FlatScene, CUDA scene, oracle scene with the fast checker BVH), cached for the session."""
    import orc
    from spt import core, scenes

    def get(name, **kw):
        key = (name, tuple(sorted(kw.items())))
        if key not in _scene_cache:
            fs = scenes.load(name, **kw)
            _scene_cache[key] = (fs, core.Scene(ctx, fs), orc.OracleScene(fs, fix_q1=True))
        return _scene_cache[key]
    return get
