import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


def pytest_report_header(config):
    import __graft_entry__ as g
    ref = g.load_oracle()
    return "parity oracle: %s" % (ref.JUDGE_KIND if ref.have_ref() else
                                   "MISSING (oracle/_ref/libref_cl.so absent: every GPU parity test will fail, by design)")


@pytest.fixture(scope="session")
def pkg():
    import __graft_entry__ as g
    return g.load_package()


@pytest.fixture(scope="session")
def ref():
    import __graft_entry__ as g
    return g.load_oracle()


@pytest.fixture(scope="session")
def scene(ref):
    return ref.render_map_scene(ROOT)


@pytest.fixture(scope="session")
def engine(pkg, scene):
    """One engine on cuda:0 holding scenes/render.map and the default image arrays."""
    e = pkg.Engine(0)
    e.upload_scene(scene.spheres, scene.planes, scene.lights)
    e.upload_images(0, scene.tex)
    e.upload_images(1, scene.sky)
    yield e
    e.close()
