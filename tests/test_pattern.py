"""The 256x4 BRIEF sampling table: both copies carry the pinned SHA-256 and, where the reference
tree is present (build container), equal the numbers in src/ORBextractor.cc:149-407."""
import hashlib
import re
import struct
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
SHA = "2164181aea6ff9ac426ca512d5130d15e1f6e3cd47b1cbdd568bbe1e55d49023"


def _nums(path):
    txt = "\n".join(l for l in path.read_text().splitlines() if not l.startswith("//"))
    return [int(v) for v in re.findall(r"-?\d+", txt)]


@pytest.mark.parametrize("rel", ["oracle/orb_pattern_oracle.inc", "orb_slam3_vio_fixes_b200/csrc/orb_pattern.inc"])
def test_pattern_checksum(rel):
    n = _nums(ROOT / rel)
    assert len(n) == 1024 and max(abs(v) for v in n) <= 13
    assert hashlib.sha256(struct.pack("<1024b", *n)).hexdigest() == SHA


def test_pattern_equals_reference_when_available():
    ref = Path("/root/reference/src/ORBextractor.cc")
    if not ref.exists():
        pytest.skip("reference tree not on this box")
    import sys
    sys.path.insert(0, str(ROOT / "tools"))
    import gen_pattern
    assert gen_pattern.read_reference_pattern() == _nums(ROOT / "oracle" / "orb_pattern_oracle.inc")
