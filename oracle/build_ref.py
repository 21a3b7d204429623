"""TEST INFRASTRUCTURE - builds the UNMODIFIED reference (PythonOT/POT) into oracle/_ref/ so that bench.py's
reference arm and the golden generators can import the real `ot` package.

    python oracle/build_ref.py [--force]

Recipe: the hot path of the reference is pure Python, but `import ot` pulls three Cython
extensions (ot.lp, ot.partial, ot.bsp).  The sources are copied from /root/reference to a scratch directory
(the reference tree is read-only and the build writes generated C++ next to the .pyx files), built there with
`setup.py build_ext --inplace`, and ONLY the resulting importable package `ot/` (python files + the three
built .so files) is placed in oracle/_ref/ot.  oracle/_ref/ is git-ignored (no reference source enters the
history) but not gpurun-ignored, so it travels to the GPU box like pot_b200/libb2ot.so.

Nothing in pot_b200/ imports this; the users are bench.py --impl reference / cpu_baseline, tests/golden/make_*.py
and tests that pin the oracle.  When /root/reference is absent (the GPU box) this script does nothing.
"""
import os
import shutil
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = "/root/reference"
OUT = os.path.join(HERE, "_ref")


def ref_available():
    return os.path.isfile(os.path.join(OUT, "ot", "__init__.py")) and os.path.isfile(os.path.join(OUT, "BUILD_OK"))


def build(force=False, jobs=8):
    if ref_available() and not force:
        return OUT
    if not os.path.isdir(os.path.join(REF_SRC, "ot")):
        return None
    tmp = tempfile.mkdtemp(prefix="potref_")
    try:
        for name in ("ot", "setup.py", "README.md", "pyproject.toml", "setup.cfg"):
            src = os.path.join(REF_SRC, name)
            if os.path.isdir(src):
                shutil.copytree(src, os.path.join(tmp, name))
            elif os.path.exists(src):
                shutil.copy(src, os.path.join(tmp, name))
        os.makedirs(os.path.join(tmp, "deps"), exist_ok=True)
        os.symlink(os.path.join(REF_SRC, "deps", "eigen"), os.path.join(tmp, "deps", "eigen"))
        subprocess.run(["chmod", "-R", "u+w", tmp], check=True)
        subprocess.run([sys.executable, "setup.py", "build_ext", "--inplace", f"-j{jobs}"], cwd=tmp, check=True,
                       stdout=subprocess.DEVNULL)
        if os.path.isdir(OUT):
            shutil.rmtree(OUT)
        os.makedirs(OUT)
        shutil.copytree(os.path.join(tmp, "ot"), os.path.join(OUT, "ot"),
                        ignore=shutil.ignore_patterns("*.cpp", "*.c", "*.pyx", "*.pxd", "*.h", "__pycache__", "build"))
        # the extensions carry ~160 MB of debug info (Eigen templates); strip it so the tree travels cheaply
        sos = [os.path.join(r, fn) for r, _, fs in os.walk(OUT) for fn in fs if fn.endswith(".so")]
        if sos and shutil.which("strip"):
            subprocess.run(["strip", "--strip-unneeded"] + sos, check=False)
        with open(os.path.join(OUT, "BUILD_OK"), "w") as f:
            f.write("built from /root/reference by oracle/build_ref.py (setup.py build_ext --inplace)\n")
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    return OUT


def import_ref():
    """Imports and returns the reference package `ot` from oracle/_ref (raises ImportError when it is not built)."""
    if not ref_available():
        raise ImportError("oracle/_ref is not built (run python oracle/build_ref.py where /root/reference exists)")
    if OUT not in sys.path:
        sys.path.insert(0, OUT)
    import ot  # noqa: E402

    if not os.path.abspath(ot.__file__).startswith(OUT):
        raise ImportError("another `ot` package shadows oracle/_ref: %s" % ot.__file__)
    return ot


if __name__ == "__main__":
    out = build(force="--force" in sys.argv)
    print("oracle/_ref:", out if out else "reference sources not present; nothing built")
