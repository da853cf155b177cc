import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

import __graft_entry__ as graft  # noqa: E402


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def pkg():
    return graft.load_pkg()


@pytest.fixture(scope="session")
def wl(pkg):
    import importlib
    return importlib.import_module(graft.PKG_NAME + ".workload")


@pytest.fixture(scope="session", params=["text", "tape"])
def hostplan(request):
    """The g++ build of the planner headers (tests/hostplan), built on demand: once walking
    the JSON text, once walking the structural index of nbrm_tape.cuh (what the warp-per-tuple
    kernels do on the GPU, with the scalar index builder in place of the warp's)."""
    import subprocess
    d = os.path.join(ROOT, "tests", "hostplan")
    subprocess.run(["make", "-C", d], check=True, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    import hostplan_binding
    name = "libhostplan.so" if request.param == "text" else "libhostplan_tape.so"
    hp = hostplan_binding.HostPlan(os.path.join(d, name))
    assert hp.lib.hostplan_tape_mode() == (request.param == "tape")
    return hp


@pytest.fixture(scope="session")
def ctx(pkg):
    """A libnbrm context on cuda:0; fails loudly (no fallback) without the library or a GPU."""
    c = pkg.Context(0)
    yield c
    c.close()
