"""CPU-side checks of the boundary: the library loads and exports every symbol include/enarf_b200.h declares."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def built():
    import __graft_entry__ as ge
    ge.build()
    from enarf_gan_b200 import _abi
    return _abi


def test_header_symbols_exported(built):
    header = open(os.path.join(ROOT, "include", "enarf_b200.h")).read()
    declared = set(re.findall(r"\b(enarf_[a-z0-9_]+)\s*\(", header))
    assert declared, "no declarations parsed"
    handle = ctypes.CDLL(built.LIB_PATH)
    for name in declared:
        assert hasattr(handle, name), f"{name} declared in the header but not exported"
    assert declared == set(built.EXPORTS), declared ^ set(built.EXPORTS)


def test_scene_struct_matches_header(built):
    # 4 ints, 3 ptr, ptr, 2 ints, ptr, int64, 6 ptr  -> 16 + 24 + 8 + 8 + 8 + 8 + 48 = 120 bytes on LP64
    assert ctypes.sizeof(built.Scene) == 120


def test_argument_errors_do_not_need_a_gpu(built):
    lib = built.lib()
    rc = lib.enarf_near_far(None, 1, 1, 0.3, 5.0, None, None)
    assert rc == 1
    assert b"enarf_near_far" in lib.enarf_last_error()
    rc = lib.enarf_fine_sample(None, 0, 48, 48, 0, None, None, 0, None, 0, None, None, None)
    assert rc == 1


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "enarf-gan_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            src = open(os.path.join(pkg, fn)).read()
            assert "enarf_oracle" not in src and "import oracle" not in src, fn
