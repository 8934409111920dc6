"""CPU checks of the drop-in boundary: the C-ABI library loads and exports every symbol include/b200_sde.h declares."""
import ctypes
import os
import re

from conftest import ROOT


def _declared():
    src = open(os.path.join(ROOT, "include", "b200_sde.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_entry_points():
    names = _declared()
    assert "b200_upfirdn2d_f32" in names and len(names) >= 10


def test_library_exports_every_declared_symbol():
    from b_pinn_kalman_filter_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        from b_pinn_kalman_filter_b200 import build
        build.build()
    lib = ctypes.CDLL(_lib.LIB_PATH)
    missing = [n for n in _declared() if not hasattr(lib, n)]
    assert not missing, missing
    assert lib.b200_abi_version() >= 1


def test_python_binding_table_matches_header():
    from b_pinn_kalman_filter_b200 import _lib
    import b_pinn_kalman_filter_b200.models.engine  # noqa: F401  registers the engine entry points
    declared = set(_declared())
    bound = set(_lib._SIGNATURES)
    assert bound <= declared, sorted(bound - declared)
    lib = _lib.lib()
    assert lib.b200_upfirdn2d_out_size(32, 4, 1, 2, 1, 1) == 16


def test_sass_is_sm100a():
    """The shipped library holds sm_100a code only (no multi-arch dispatch)."""
    import subprocess
    from b_pinn_kalman_filter_b200 import _lib
    out = subprocess.run(["cuobjdump", "-lelf", _lib.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_\d+a?", out))
    assert archs == {"sm_100a"}, archs
