import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "dune-ddm_b200"))
sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def small3d():
    """3-D P1 Poisson, islands coefficient, 8 subdomains, overlap 2, GenEO (nev 6)."""
    from ddm_b200.setup import problem

    return problem.unitcube_p1(12, 8, overlap=2, nev=6, workers=1)


@pytest.fixture(scope="session")
def small2d():
    """2-D P1 Poisson (config C1 shape at reduced size): 16 subdomains, overlap 2, GenEO (nev 8)."""
    from ddm_b200.setup import problem

    return problem.unitsquare_p1(40, 16, overlap=2, nev=8, workers=1)


@pytest.fixture(scope="session")
def cd_dg():
    """Config C4 at reduced size: Q1-DG convection-diffusion, 48^2 cells, 6 subdomains, overlap 1, LU factors,
    POU x {1, x, y, xy} coarse space (non-symmetric A and A0)."""
    from ddm_b200.setup import problem

    return problem.convdiff_dg(48, 6, workers=1)


@pytest.fixture(scope="session")
def elast():
    """Config C3 at reduced size: P1 elasticity on the 10 x 1 x 1.5 box (30 x 3 x 4 cells), steel bars in rubber, clamped
    at x = 0, 6 subdomains, overlap 1, GenEO nev 8; node-major DOFs (3x3 block operator)."""
    from ddm_b200.setup import problem

    return problem.elasticity_box(3, 6, nev=8, workers=1)


def pytest_collection_modifyitems(config, items):
    """GPU tests get a hard time limit (pytest-timeout, thread method: the process is ended even when the main thread sits in a
    CUDA synchronize), so that a kernel that never finishes costs one test run and not the GPU box."""
    if not config.pluginmanager.hasplugin("timeout"):
        return
    for item in items:
        if item.get_closest_marker("gpu") and not item.get_closest_marker("timeout"):
            item.add_marker(pytest.mark.timeout(900, method="thread"))


@pytest.fixture(autouse=True)
def _gpu_tests_run_on_the_real_library(request):
    """Every `-m gpu` test must exercise libddm_b200.so itself: the dry-run stand-in (tests/dryrun_standin.py, an opt-in
    proof-reading plugin that answers the C ABI from the oracle) reports version b"dryrun" and must never count as GPU coverage."""
    if request.node.get_closest_marker("gpu") and not request.config.pluginmanager.hasplugin("dryrun_standin"):
        from ddm_b200 import lib

        assert b"dryrun" not in lib.load().ddm_version(), "GPU test running on the dry-run stand-in"
    yield
