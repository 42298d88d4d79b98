"""CPU-side checks of the C-ABI boundary: the library builds, loads, and exports exactly what
include/evoxmo.h declares (no compute calls - those need a GPU and live in test_gpu_*.py)."""
import ctypes
import re
import subprocess
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent


@pytest.fixture(scope="module")
def so():
    from evox_b200.build import build
    return build()


def _declared():
    text = (ROOT / "include" / "evoxmo.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(evx_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_exported(so):
    names = _declared()
    assert len(names) >= 28
    L = ctypes.CDLL(str(so))
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing


def test_binding_matches_header(so):
    from evox_b200._lib import exported_symbols, lib
    lib()
    assert sorted(exported_symbols()) == _declared()


def test_no_cxx_leak_in_abi(so):
    out = subprocess.run(["nm", "-D", "--defined-only", str(so)], capture_output=True, text=True, check=True).stdout
    evx = [l.split()[-1] for l in out.splitlines() if " T " in l and "evx_" in l.split()[-1]]
    assert all(not s.startswith("_Z") for s in evx if s.startswith("evx_"))
    assert "evx_nds_f32" in evx


def test_sm100a_only(so):
    out = subprocess.run(["cuobjdump", "-lelf", str(so)], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs


def test_host_only_entry_points(so):
    from evox_b200._lib import lib
    from oracle import oracle_np as O
    L = lib()
    assert L.evx_version() >= 100
    for d, m in [(300, 3), (120, 2), (500, 5), (77, 3), (1000, 10), (7, 3)]:
        sub = (ctypes.c_int32 * m)()
        off = (ctypes.c_int32 * (m + 1))()
        assert L.evx_lsmop_groups(d, m, sub, off) == 0
        s, o = O.lsmop_groups(d, m)                 # lsmop.py:46-54
        assert tuple(sub) == s and tuple(off) == o
    # workspace sizing: packed triangular bit matrix ~ n^2/16 bytes
    b = L.evx_nds_workspace_bytes(1_000_000, 3)
    assert 62e9 < b < 64e9
    assert L.evx_nds_workspace_bytes(10, 0) == 0 and L.evx_nds_workspace_bytes(10, 33) == 0
    w8 = sum(L.evx_nds_shard_workspace_bytes(1_000_000, 3, 8, r) for r in range(8))
    assert w8 < b * 1.05
    # argument errors are reported, not raised across the ABI
    assert L.evx_nds_f32(None, 5, 3, None, None, 0, None) != 0
    assert b"null" in L.evx_last_error()


def test_product_never_imports_oracle():
    for p in (ROOT / "evox_b200").rglob("*.py"):
        assert "oracle" not in p.read_text(), p
