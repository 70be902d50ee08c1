"""The C-ABI library loads and exports every symbol include/gffx_b200.h declares (no compute calls)."""
import ctypes as C
import os
import re
import subprocess

import pytest

from tests.conftest import ROOT


def header_symbols():
    text = open(os.path.join(ROOT, "include", "gffx_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(gffx_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_exported(gffx):
    from gffx_b200 import _lib

    syms = header_symbols()
    assert len(syms) >= 40
    out = subprocess.check_output(["nm", "-D", "--defined-only", _lib.LIB_PATH]).decode()
    exported = set(l.split()[-1] for l in out.splitlines() if l.strip())
    missing = [s for s in syms if s not in exported]
    assert not missing, missing
    for s in syms:
        assert hasattr(_lib.lib, s)
    # the ctypes table covers the whole header, nothing more
    assert sorted(_lib.SIGNATURES) == syms


def test_no_oracle_or_torch_in_product_library(gffx):
    """The product library must not link the oracle, nor torch (the boundary is a plain C ABI)."""
    from gffx_b200 import _lib

    out = subprocess.check_output(["ldd", _lib.LIB_PATH]).decode()
    assert "oracle" not in out and "torch" not in out
    for dirpath, _, files in os.walk(os.path.join(ROOT, "gffx_b200")):
        for f in files:
            if f.endswith((".py", ".cpp", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f), errors="replace").read()
                assert "gffx_oracle" not in src and "liboracle" not in src, f


def test_version_and_errors(gffx):
    from gffx_b200 import _lib

    assert b"sm_100a" in _lib.lib.gffx_version()
    h = C.c_void_p()
    rc = _lib.lib.gffx_index_open(b"/nonexistent/x.gff", -1, C.byref(h))
    assert rc == -2 and "SQS" in _lib.last_error()
    rc = _lib.lib.gffx_index_open(None, -1, C.byref(h))
    assert rc == -1


def test_compute_fails_loudly_without_gpu(gffx, kat_gff):
    """No CPU fallback: with no usable device the compute entry points return GFFX_E_CUDA."""
    import numpy as np

    from tests.conftest import has_cuda

    idx = gffx.TreeIndexData.load_tree_index(kat_gff, device=-1)
    z = np.zeros(1, dtype=np.uint32)
    with pytest.raises(gffx.GffxError) as ei:
        idx.query_features((z, z, z + 5))
    assert ei.value.code == -4 and "no CPU fallback" in ei.value.msg
    if not has_cuda():
        with pytest.raises(gffx.GffxError) as ei:
            gffx.TreeIndexData.load_tree_index(kat_gff, device=0)
        assert ei.value.code == -4


def test_rust_crate_sources_match_the_makefile():
    """cuda/build.rs (the crate a GFFx maintainer adds) compiles the very source lists of the Makefile, and
    cuda/src/lib.rs declares only functions the header has."""
    mk = open(os.path.join(ROOT, "Makefile")).read()
    cu = re.findall(r"\$\(SRC\)/(\S+\.cu)\b", re.search(r"^CU_SRCS := (.*)$", mk, re.M).group(1))
    cpp = re.findall(r"\$\(SRC\)/(\S+\.cpp)\b", re.search(r"^CPP_SRCS := (.*)$", mk, re.M).group(1))
    rs = open(os.path.join(ROOT, "cuda", "build.rs")).read()
    got_cu = re.findall(r'"([^"]+\.cu)"', re.search(r"let cu = \[(.*?)\];", rs, re.S).group(1))
    got_cpp = re.findall(r'"([^"]+\.cpp)"', re.search(r"let cpp = \[(.*?)\];", rs, re.S).group(1))
    assert sorted(got_cu) == sorted(cu) and sorted(got_cpp) == sorted(cpp)
    for f in cu + cpp:
        assert os.path.exists(os.path.join(ROOT, "gffx_b200", "csrc", f)), f
    assert "sm_100a" in rs
    lib_rs = open(os.path.join(ROOT, "cuda", "src", "lib.rs")).read()
    declared = set(re.findall(r"\bfn (gffx_[a-z0-9_]+)\s*\(", lib_rs))
    assert len(declared) >= 20 and declared <= set(header_symbols()), sorted(declared - set(header_symbols()))
    assert "window.cu" not in open(os.path.join(ROOT, "INTEGRATION.md")).read()
