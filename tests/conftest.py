import os
import sys
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def pkg():
    import __graft_entry__ as g
    return g.load_package()


@pytest.fixture(scope="session")
def oracle():
    import __graft_entry__ as g
    O = g.load_oracle()
    O.build()
    return O


@pytest.fixture(scope="session")
def window_small(pkg):
    return pkg.synthetic.make_window(seed=1, w=320, h=240, nF=4, n_points=500)


@pytest.fixture(scope="session")
def window_vga(pkg):
    return pkg.synthetic.make_window(seed=0)
