import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle as O
    O.build()
    return O


@pytest.fixture(scope="session")
def synth():
    from vslam_b200 import synth as S
    S.build()
    return S


@pytest.fixture(scope="session")
def capi():
    from vslam_b200 import build, capi as A
    if not os.path.exists(A.LIB_PATH):
        build.build_cuda()
    return A


@pytest.fixture(params=["chi2-reference", "chi2-weighted"])
def chi2_mode(request, oracle, capi):
    """Runs a test under both definitions of a view's chi2 (vslam_b200.h): make_config() of the oracle and of the
    binding default to the chosen one for the duration of the test."""
    mode = 0 if request.param == "chi2-reference" else 1
    old = oracle.DEFAULT_CHI2_MODE, capi.DEFAULT_CHI2_MODE
    oracle.DEFAULT_CHI2_MODE = capi.DEFAULT_CHI2_MODE = mode
    yield mode
    oracle.DEFAULT_CHI2_MODE, capi.DEFAULT_CHI2_MODE = old


@pytest.fixture()
def ctx(capi, request):
    if "level_kernel" in request.fixturenames:  # the kernel choice reaches make_config through the environment
        request.getfixturevalue("level_kernel")
    if "chi2_mode" in request.fixturenames:
        request.getfixturevalue("chi2_mode")
    c = capi.Context(capi.make_config())
    yield c
    c.close()
