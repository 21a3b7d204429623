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
