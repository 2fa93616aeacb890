"""The C-ABI library loads and exports every symbol include/gnnuds.h declares (no compute calls)."""
import ctypes
import os
import re
import subprocess

import pytest

import cases
from gnn_uds_b200 import _lib, build

HDR = os.path.join(cases.ROOT, "include", "gnnuds.h")


@pytest.fixture(scope="module")
def lib_path():
    return build.build()


def _declared():
    src = open(HDR).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(gnnuds_[a-z_]+)\s*\(", src)))


def test_header_symbols_exported(lib_path):
    out = subprocess.run(["nm", "-D", "--defined-only", lib_path], capture_output=True, text=True, check=True).stdout
    exported = set(re.findall(r" T (gnnuds_[a-z_]+)", out))
    declared = _declared()
    assert declared, "no declarations parsed"
    assert not [s for s in declared if s not in exported]
    assert sorted(_lib.SYMBOLS) == declared          # the ctypes shim binds exactly the header


def test_integration_doc_names_every_entry_point():
    """INTEGRATION.md's entry-point table (what each C symbol replaces in the reference) covers the whole header; a
    `gnnuds_a/_b` shorthand counts for both names."""
    doc = open(os.path.join(cases.ROOT, "INTEGRATION.md")).read()
    missing = [s for s in _declared() if s not in doc and not re.search(r"/_%s\b" % re.escape(s.rsplit("_", 1)[-1]), doc)]
    assert not missing, missing


def test_ctypes_load_and_struct_sizes(lib_path):
    lib = _lib.load()
    assert ctypes.sizeof(_lib.Config) == 30 * 4
    for s in _lib.SYMBOLS:
        assert hasattr(lib, s)


def test_sass_is_sm100a(lib_path):
    out = subprocess.run(["cuobjdump", "-lelf", lib_path], capture_output=True, text=True).stdout
    assert "sm_100a" in out


def test_no_cpu_fallback(lib_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(_lib.GnnUdsError, match="no CUDA device"):
        cases.make_product("c1_ast_gcn_gru")


def test_product_does_not_import_oracle():
    pkg = os.path.join(cases.ROOT, "gnn_uds_b200")
    for f in os.listdir(pkg):
        if f.endswith(".py"):
            src = open(os.path.join(pkg, f)).read()
            assert not re.search(r"^\s*(from|import)\s+oracle", src, flags=re.M), f
