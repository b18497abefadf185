import os
import sys

import pytest

# the multi-rank emulation tests run up to 8 scenes (16 streams) with spinning mailbox waits on ONE device: enough hardware
# queues that no stream is serialised behind another scene's wait kernel (must be set before the CUDA context exists)
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

WORLDS = ["world1", "world2", "world4", "world8", "world16", "world8_stress"]
N_INSTANCES = {"world1": 2, "world2": 14, "world4": 62, "world8": 380, "world16": 1452, "world8_stress": 570,
               "config": 1452}


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")


@pytest.fixture(scope="session")
def pkg():
    import rtb200_loader
    p = rtb200_loader.load()
    if not os.path.exists(p.HOST_LIB_PATH) or not os.path.exists(p.CUDA_LIB_PATH):
        import __graft_entry__
        __graft_entry__.build()
    return p


def scene_path(name):
    return os.path.join(ROOT, "scenes", name + ".json")


def golden(name):
    import json
    import numpy as np
    meta = json.load(open(os.path.join(ROOT, "tests", "golden", name + ".json")))
    data = np.load(os.path.join(ROOT, "tests", "golden", name + ".npz"))
    return meta, data


@pytest.fixture(scope="session")
def gpu_ok(pkg):
    n = pkg.C.c_int(0)
    rc = pkg.cuda_lib().rt_device_count(pkg.C.byref(n))
    if rc != 0 or n.value <= 0:
        pytest.fail("no CUDA device visible: -m gpu tests must run on the GPU box (there is no CPU fallback)")
    return n.value
