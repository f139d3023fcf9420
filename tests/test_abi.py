"""CPU-side checks of the C-ABI boundary: the library loads without a GPU and exports exactly what
include/grl_b200.h declares, with the arity the ctypes binding uses.  No compute calls."""
import ctypes
import os
import re
import subprocess

import pytest

from conftest import ROOT

HEADER = os.path.join(ROOT, "include", "grl_b200.h")


def _declared():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    decls = {}
    for m in re.finditer(r"\b(?:int|size_t|const char\*)\s+(grl_\w+)\s*\(([^;]*?)\)\s*;", src, flags=re.S):
        args = m.group(2).strip()
        n = 0 if args in ("", "void") else len([a for a in args.split(",") if a.strip()])
        decls[m.group(1)] = n
    return decls


@pytest.fixture(scope="module")
def lib_path():
    from graph_representation_learning_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        import __graft_entry__
        __graft_entry__.build()
    return _lib.LIB_PATH


def test_header_declares_entry_points():
    d = _declared()
    assert len(d) >= 15
    for name in ("grl_csr_count", "grl_csr_fill", "grl_csr_transpose", "grl_spmm_fwd", "grl_spmm_bwd",
                 "grl_dropedge_weights", "grl_bias_act_dropout_fwd", "grl_last_error"):
        assert name in d


def test_library_exports_every_declared_symbol(lib_path):
    lib = ctypes.CDLL(lib_path)
    for name in _declared():
        assert hasattr(lib, name), f"{name} declared in grl_b200.h but not exported"
    out = subprocess.run(["nm", "-D", "--defined-only", lib_path], capture_output=True, text=True).stdout
    exported = {ln.split()[-1] for ln in out.splitlines() if ln.strip()}
    extra = {s for s in exported if not s.startswith("grl_")}
    assert not extra, f"non-ABI symbols leak from the library: {sorted(extra)[:5]}"
    assert {s for s in exported if s.startswith("grl_")} == set(_declared())


def test_ctypes_binding_matches_header(lib_path):
    from graph_representation_learning_b200 import _lib
    d = _declared()
    assert set(_lib.SIGNATURES) == set(d)
    for name, (_, args) in _lib.SIGNATURES.items():
        assert len(args) == d[name], f"{name}: binding has {len(args)} args, header {d[name]}"
    lib = _lib.load()
    assert lib.grl_abi_version() == 1


def test_argument_errors_are_reported_without_gpu(lib_path):
    """Argument validation happens before any CUDA call, so it is testable here."""
    from graph_representation_learning_b200 import _lib
    lib = _lib.load()
    rc = lib.grl_spmm_fwd(0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 0)
    assert rc == -1 and b"null pointer" in lib.grl_last_error()
    rc = lib.grl_dropedge_weights(1, 0, 4, 1.5, 0, 0, 1, 0)
    assert rc == -1 and b"outside" in lib.grl_last_error()
    with pytest.raises(_lib.GrlError):
        _lib.check(rc, "grl_dropedge_weights")


def test_sass_is_sm100a_only(lib_path):
    out = subprocess.run(["cuobjdump", "-lelf", lib_path], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_\d+a?", out))
    assert archs == {"sm_100a"}, archs


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "graph_representation_learning_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(".py"):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f
