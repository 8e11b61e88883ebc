"""CPU: libctrhot.so builds for sm_100a, loads, and exports every symbol include/ctrhot.h declares.
No compute calls here (no GPU); argument validation paths that return before any CUDA call are exercised."""
import ctypes
import os
import re
import subprocess

import pytest

from recommender_system_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_header_parses_every_declaration():
    src = open(os.path.join(ROOT, "include", "ctrhot.h")).read()
    src = re.sub(r"/\*.*?\*/", " ", src, flags=re.S)
    names = set(re.findall(r"\b(ctrhot_\w+)\s*\(", src))
    assert names == set(_lib.SIGNATURES), names ^ set(_lib.SIGNATURES)
    assert len(names) >= 30


def test_library_exports_every_declared_symbol(lib):
    for name in _lib.SIGNATURES:
        assert hasattr(lib, name), name


def test_no_undeclared_exports(lib):
    out = subprocess.run(["nm", "-D", "--defined-only", _lib.LIB_PATH], capture_output=True, text=True).stdout
    exported = {l.split()[-1] for l in out.splitlines() if " T " in l and l.split()[-1].startswith("ctrhot_")}
    assert exported == set(_lib.SIGNATURES), exported ^ set(_lib.SIGNATURES)


def test_sass_is_sm100a_with_tcgen05_and_bulk_copy(lib):
    """The tower kernel must contain the Blackwell tensor path (UTC*MMA / LDTM) and the TMA-engine bulk copy."""
    out = subprocess.run(["cuobjdump", "-sass", _lib.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in out
    assert re.search(r"UTC\w*MMA", out), "no tcgen05.mma in SASS"
    assert "LDTM" in out, "no tcgen05.ld in SASS"
    assert "UBLKCP" in out, "no cp.async.bulk in SASS"
    assert not re.search(r"(?<![A-Z])HMMA", out), "legacy mma.sync path present"


def test_version_and_status_strings(lib):
    assert lib.ctrhot_version() == 100
    assert lib.ctrhot_status_str(0) == b"ok"
    assert b"range" in lib.ctrhot_status_str(-4)


def test_bad_arguments_return_earg_without_touching_cuda(lib):
    assert lib.ctrhot_gather_f32(0, 0, 26, 16, 0, 0, 26, 8, 0, 0, 0, 0, 416, 0, 0) == _lib.EARG
    assert lib.ctrhot_fm_dense_f32(0, 0, 0, 0, 0, 0, 0, 8, 0, 0, 0) == _lib.EARG
    assert lib.ctrhot_cross_f32(0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0) == _lib.EARG
    assert lib.ctrhot_mlp_f32(0, 0, 0, None, 0, 0, 0, 0, 0, 0) == _lib.EARG
    assert lib.ctrhot_mlp_pack_bytes(0, 4) == _lib.EARG
    with pytest.raises(ValueError):
        _lib.check(_lib.EARG, "x")
    with pytest.raises(IndexError):
        _lib.check(_lib.EID, "x")


def test_pack_sizes(lib):
    # 429 -> 256: 14 K-chunks x 256 rows x 128 B
    assert lib.ctrhot_mlp_pack_bytes(429, 256) == 14 * 256 * 128
    # 400 wide splits into two 208-column tiles
    assert lib.ctrhot_mlp_pack_bytes(400, 400) == 2 * 13 * 208 * 128
    assert lib.ctrhot_fm_planes_bytes(429, 10) == 12 * 448 * 4


def test_dense_desc_layout_matches_header():
    assert ctypes.sizeof(_lib.DenseDesc) == 4 * 8 + 4 * 4 + 4 + 4
