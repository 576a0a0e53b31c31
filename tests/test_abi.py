"""The C-ABI library loads and exports exactly the symbols include/b200infer.h declares (no compute calls: no GPU here)."""
import ctypes
import os
import re
import subprocess

import rdb_b200
from rdb_b200 import lib as L

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, "include", "b200infer.h")).read()
    return set(re.findall(r"B200_API\s+[\w\s\*]+?\b(b200_\w+)\s*\(", src))


def test_every_declared_symbol_is_exported_and_bound():
    lib = L.load()
    declared = header_symbols()
    assert len(declared) >= 20
    for name in declared:
        assert hasattr(lib, name), "missing export %s" % name
    assert declared == set(L.SIGNATURES), (declared ^ set(L.SIGNATURES))


def test_no_undeclared_b200_exports():
    out = subprocess.run(["nm", "-D", "--defined-only", L.lib_path()], capture_output=True, text=True).stdout
    exported = {l.split()[-1] for l in out.splitlines() if " T " in l and l.split()[-1].startswith("b200_")}
    assert exported == header_symbols()


def test_abi_version_and_error_plumbing_without_gpu():
    lib = L.load()
    assert lib.b200_abi_version() == 1
    h = ctypes.c_void_p()
    rc = lib.b200_engine_create(b"resnet50", b"not a blob", 10, 4, 0, ctypes.byref(h))
    assert rc == -3 and b"blob" in lib.b200_last_error()
    assert not h.value


def test_library_has_blackwell_native_sass():
    """tcgen05.mma -> UTCHMMA, TMA -> UTMALDG/UTMASTG, tcgen05.ld -> LDTM (B200_PROFILING.md)."""
    r = subprocess.run(["cuobjdump", "-sass", L.lib_path()], capture_output=True, text=True)
    if r.returncode != 0:
        import pytest
        pytest.skip("cuobjdump unavailable")
    for mnemonic in ("UTCHMMA", "UTMALDG", "UTMASTG", "LDTM"):
        assert mnemonic in r.stdout, mnemonic
