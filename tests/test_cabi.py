"""CPU-side checks of the drop-in boundary: the C-ABI library loads (no GPU needed) and exports every
symbol include/b2ot.h declares; error paths that do not touch the device behave."""
import ctypes
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    so = os.path.join(ROOT, "pot_b200", "libb2ot.so")
    if not os.path.exists(so):
        subprocess.run(["make", "-C", os.path.join(ROOT, "pot_b200", "csrc")], check=True)
    from pot_b200 import _lib
    return _lib.load()


def declared_symbols():
    out = []
    for fn in sorted(os.listdir(os.path.join(ROOT, "include"))):
        if fn.endswith(".h"):
            src = open(os.path.join(ROOT, "include", fn)).read()
            src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
            out += re.findall(r"\b(b2ot_[a-z0-9_]+)\s*\(", src)
    return sorted(set(s for s in out if not s.endswith("_fn")))


def test_every_declared_symbol_is_exported(lib):
    syms = declared_symbols()
    assert len(syms) >= 15
    from pot_b200 import _lib
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/*.h but not exported by libb2ot.so"
        assert s in _lib.PROTOTYPES, f"{s} has no ctypes prototype in pot_b200/_lib.py"


def test_version_and_error_reporting(lib):
    assert lib.b2ot_version() >= 100
    from pot_b200 import _lib
    ok, info = _lib.build_matches_sources()
    assert ok, f"stale library: {info}, sources hash {_lib.source_hash()}"
    h = ctypes.c_void_p()
    rc = lib.b2ot_dense_create(ctypes.byref(h), 7, 0, 4, 4, 4, None, None, None, None, None, 0, 10, 10, 0.0, 0, None)
    assert rc < 0
    assert b"bad mode" in lib.b2ot_last_error()
    rc = lib.b2ot_kexp(None, None, 4, 4, 4, 4, 1.0, 9, None)
    assert rc < 0 and b"dtype" in lib.b2ot_last_error()
    assert lib.b2ot_dense_workspace_bytes(1000, 1000, 1, 1000, 10) > 3 * 1000 * 8


def test_no_cpu_fallback():
    import numpy as np
    import torch
    import pot_b200 as ot
    M = torch.eye(3)
    with pytest.raises(ValueError, match="no CPU fallback"):
        ot.sinkhorn([], [], M, 1.0)
    with pytest.raises(TypeError):
        ot.sinkhorn([], [], np.eye(3), 1.0)
    with pytest.raises(ValueError, match="Unknown method"):
        ot.sinkhorn([], [], M, 1.0, method="bogus")


def test_product_does_not_import_oracle():
    """The oracle is test infrastructure: nothing under pot_b200/ may import, call or even mention it."""
    pkg = os.path.join(ROOT, "pot_b200")
    for dp, _, fns in os.walk(pkg):
        for fn in fns:
            if fn.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dp, fn)).read().lower()
                assert "oracle" not in text, f"{fn} mentions the oracle: the product path must not depend on it"


def test_argument_checks_need_no_gpu(lib):
    """Entry points validate their arguments before any CUDA call: bad calls return -1 with a message, on any machine."""
    import ctypes as C

    from pot_b200 import _lib

    L = _lib.load()
    dummy = C.c_void_p(4096)  # never dereferenced: the checks come first
    maxc = 16384  # B2OT_COLSUM_STATIC_MAX_COLS
    assert L.b2ot_colsum_static(_lib.F32, dummy, 2, maxc + 1, maxc + 1, dummy, None) == -1
    msg = L.b2ot_last_error().decode()
    assert "exceeds the static scratch" in msg and str(maxc) in msg
    assert L.b2ot_colsum_static(7, dummy, 2, 8, 8, dummy, None) == -1
    assert "bad dtype" in L.b2ot_last_error().decode()
    assert L.b2ot_colsum_static(_lib.F32, dummy, 2, 8, 4, dummy, None) == -1  # ld < m
    d2 = C.c_void_p(8192)  # u_new / v_new must differ from u_prev / v_prev
    args = (dummy, 4, 5, 5, 2, dummy, dummy, dummy, dummy, d2, d2, 0, 10, 0.0, 1.0, 1.0, dummy, dummy, 4, dummy, None)
    assert L.b2ot_multi_scaling_step(_lib.F64, 9, *args) == -1
    assert "bad mode" in L.b2ot_last_error().decode()
    bad_fi = args[:14] + (1.5,) + args[15:]
    assert L.b2ot_multi_scaling_step(_lib.F64, _lib.MULTI_UNBALANCED, *bad_fi) == -1
    assert "exponents" in L.b2ot_last_error().decode()
    assert L.b2ot_multi_absorb(_lib.F64, 0, 5, 2, dummy, dummy, dummy, dummy, None) == -1
    assert L.b2ot_multi_scratch_bytes(100, 1000, 3) == (32 * 3 * 1000 + 2 * 4 * 3) * 8


def test_header_is_plain_c_and_links_from_c(tmp_path):
    """include/b2ot.h compiles as C and a C program links against libb2ot.so (extern "C", no C++ or torch types)."""
    import shutil
    import subprocess

    if shutil.which("gcc") is None:
        pytest.skip("no gcc")
    src = tmp_path / "use_b2ot.c"
    src.write_text(
        '#include <stdio.h>\n#include <string.h>\n#include "b2ot.h"\n'
        "int main(void) {\n"
        '  printf("%d %s\\n", b2ot_version(), b2ot_build_info());\n'
        "  int rc = b2ot_colsum_static(B2OT_F32, (const void*)4096, 2, B2OT_COLSUM_STATIC_MAX_COLS + 1,\n"
        "                              B2OT_COLSUM_STATIC_MAX_COLS + 1, (void*)4096, 0);\n"
        '  return (rc == -1 && strstr(b2ot_last_error(), "exceeds") != 0) ? 0 : 1;\n'
        "}\n")
    exe = tmp_path / "use_b2ot"
    libdir = os.path.join(ROOT, "pot_b200")
    r = subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe),
                        "-L", libdir, "-lb2ot", f"-Wl,-rpath,{libdir}"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    r = subprocess.run([str(exe)], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "src=" in r.stdout
