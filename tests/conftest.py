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


@pytest.fixture(scope="session")
def window_large(pkg):
    """one BASELINE configs[3] size under the driver-run tests: 1280x1024, ~18 k active points (level-0 cloud ~55 k points,
    which selects the multi-cluster tracker path), 4 key frames, ~54 k residuals"""
    return pkg.synthetic.make_window(seed=0, w=1280, h=1024, nF=4, n_points=18000, kf_step_t=0.04, kf_step_deg=1.0)
