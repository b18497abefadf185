"""CPU: the C-ABI libraries load, export every symbol the headers declare, and fail loudly (no CPU
fallback) when there is no GPU."""
import ctypes as C
import os
import re

import pytest

from conftest import ROOT, scene_path


def declared(header):
    src = open(os.path.join(ROOT, "include", header)).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b((?:rt|rth)_[a-z0-9_]+)\s*\(", src)))


def test_engine_exports_every_declared_symbol(pkg):
    lib = pkg.cuda_lib()
    names = declared("rtb200.h")
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), n
    assert b"sm_100a" in lib.rt_version()


def test_host_exports_every_declared_symbol(pkg):
    lib = pkg.host_lib()
    for n in declared("rtb200_host.h"):
        assert hasattr(lib, n), n


def test_no_cpu_fallback(pkg):
    """Without a CUDA device every compute entry point returns an error; nothing renders on the CPU."""
    lib = pkg.cuda_lib()
    n = C.c_int(0)
    lib.rt_device_count(C.byref(n))
    if n.value > 0:
        pytest.skip("a GPU is present")
    w = pkg.World(scene_path("world1"))
    with pytest.raises(RuntimeError):
        pkg.Scene(w)
    assert lib.rt_init(0) != 0
    assert len(lib.rt_last_error_string()) > 0


def test_product_does_not_reference_oracle():
    """Nothing under the package may import, link or mention the oracle library."""
    pkg_dir = os.path.join(ROOT, "gpu-ray-tracer_b200")
    for d, _, files in os.walk(pkg_dir):
        for f in files:
            if f.endswith((".cu", ".cuh", ".h", ".cc", ".py")):
                for line in open(os.path.join(d, f), errors="ignore"):
                    if re.search(r"#\s*include|import |CDLL|dlopen", line):
                        assert not re.search(r"oracle|lbvh_host", line), (f, line)
    out = os.popen("ldd " + os.path.join(pkg_dir, "librtb200.so")).read()
    assert "oracle" not in out
