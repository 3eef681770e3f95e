"""The C-ABI library loads and exports every symbol include/tcb200.h declares (no compute, no GPU)."""
import ctypes
import os
import re

from conftest import ROOT
from triton_client_b200 import _lib


def _declared():
    text = open(os.path.join(ROOT, "include", "tcb200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(tcb_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_entry_points():
    names = _declared()
    assert "tcb_create" in names and "tcb_preprocess_image" in names and len(names) >= 8


def test_library_exports_every_declared_symbol():
    assert os.path.exists(_lib.LIB_PATH), "libtcb200.so not built (python -c 'import __graft_entry__ as g; g.build()')"
    l = ctypes.CDLL(_lib.LIB_PATH)
    missing = [n for n in _declared() if not hasattr(l, n)]
    assert not missing, f"declared in tcb200.h but not exported: {missing}"


def test_binding_table_matches_header():
    assert sorted(_lib.SIGNATURES) == _declared()


def test_version_and_error_without_gpu():
    l = _lib.lib()
    assert b"sm_100a" in l.tcb_version()
    # a null context is an argument error, reported through the thread-local message, never a crash
    assert l.tcb_workspace_bytes(None) == 0
    assert l.tcb_destroy(None) == 0


def test_no_oracle_import_in_product():
    """The product package must never import the oracle (it is test infrastructure)."""
    pkg = os.path.join(ROOT, "triton_client_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(".py"):
                src = open(os.path.join(dp, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f"{f} imports oracle"
