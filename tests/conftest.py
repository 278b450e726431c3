import ctypes as C
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle_mod():
    import oracle
    oracle.build()
    return oracle


@pytest.fixture(scope="session")
def harness():
    """CPU build of the product's host/device-shared headers (tests/cpu_harness)."""
    src = ROOT / "tests" / "cpu_harness" / "harness.cpp"
    so = ROOT / "tests" / "cpu_harness" / "libharness.so"
    csrc = ROOT / "orb_slam3_vio_fixes_b200" / "csrc"
    deps = [src] + list(csrc.glob("*.h"))
    if not so.exists() or any(d.stat().st_mtime > so.stat().st_mtime for d in deps):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-ffp-contract=off", "-mfma",
                               f"-I{csrc}", "-o", str(so), str(src)])
    H = C.CDLL(str(so))
    H.h_sincos_mismatches.restype = C.c_long
    H.h_sincos_mismatches.argtypes = [C.c_uint32, C.c_uint32, C.c_uint32]
    H.h_logf_mismatches.restype = C.c_long
    H.h_logf_mismatches.argtypes = [C.c_uint32, C.c_uint32, C.c_uint32]
    return H


def has_cuda():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False
