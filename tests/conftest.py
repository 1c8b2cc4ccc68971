import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")
    # CPU-side libraries are cheap to (re)build; the CUDA library is built by __graft_entry__.build()
    for d in ("oracle", os.path.join("pathtracer_b200", "host")):
        subprocess.run(["make", "-C", os.path.join(ROOT, d)], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)


def golden(name):
    return np.load(os.path.join(GOLDEN, name))


@pytest.fixture(scope="session")
def worlds():
    """HostWorld objects for worlds 1..5 (scene seed 7), built once."""
    from pathtracer_b200 import host
    cache = {}

    def get(w):
        if w not in cache:
            cache[w] = host.HostWorld(w, scene_seed=7)
        return cache[w]
    return get


@pytest.fixture(scope="session")
def core():
    """The CUDA render core on device 0 (gpu tests only)."""
    from pathtracer_b200 import render
    rc = render.RenderCore(0)
    yield rc
    rc.close()


@pytest.fixture(scope="session")
def ref():
    """The reference itself (oracle/_ref), when it has been built in this checkout."""
    from oracle import ref_harness
    if not ref_harness.available():
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    return ref_harness.RefHarness()
