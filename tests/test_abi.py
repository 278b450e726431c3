"""The C-ABI library loads on a CPU-only box and exports every symbol include/orbx.h declares;
compute entry points fail loudly without a device (no CPU fallback)."""
import ctypes as C
import re
import subprocess
from pathlib import Path

import numpy as np
import pytest

from orb_slam3_vio_fixes_b200 import _lib
from orb_slam3_vio_fixes_b200 import build as B

ROOT = Path(__file__).resolve().parents[1]


def _declared():
    txt = (ROOT / "include" / "orbx.h").read_text()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(orb[xm]_[a-z0-9_]+)\s*\(", txt)))


def test_header_symbols_exported():
    so = B.build()
    L = C.CDLL(str(so))
    decl = _declared()
    assert len(decl) >= 20
    for name in decl:
        assert hasattr(L, name), f"{name} declared in include/orbx.h but not exported"
    assert sorted(_lib.SYMBOLS) == decl, "ctypes binding list out of date"
    nm = subprocess.check_output(["nm", "-D", "--defined-only", str(so)], text=True)
    exported = sorted(set(re.findall(r" T (orb[xm]_\w+)", nm)))
    assert exported == decl, "library exports symbols the header does not declare (or vice versa)"


def test_struct_layouts():
    assert _lib.KP_DTYPE.itemsize == 28 and _lib.RES_DTYPE.itemsize == 12
    assert C.sizeof(_lib.OrbxConfig) == 40
    assert _lib.lib().orbx_version().startswith(b"orbx")


def test_sass_is_sm100a_only():
    so = B.build()
    out = subprocess.check_output(["cuobjdump", "-lelf", str(so)], text=True)
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs


def test_no_device_fails_loudly():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from orb_slam3_vio_fixes_b200.extractor import ORBextractor
    with pytest.raises(_lib.OrbxError) as e:
        ORBextractor(500, 1.2, 8, 20, 7)(np.zeros((300, 400), np.uint8))
    assert e.value.code == -6


def test_product_never_imports_oracle():
    for p in (ROOT / "orb_slam3_vio_fixes_b200").rglob("*"):
        if p.suffix in (".py", ".cu", ".cuh", ".h") and p.is_file():
            txt = p.read_text()
            assert not re.search(r"import\s+oracle|from\s+oracle|#include[^\n]*oracle|liborb_oracle|oracle/", txt), p
