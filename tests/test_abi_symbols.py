"""CPU-side checks of the drop-in boundary: libfimexcu.so loads without a GPU, exports every entry point that
include/fimexcu.h declares, has no dependency on oracle/ and fails loudly (no CPU fallback) without a device."""
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "fimexcu.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(fimexcu_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    import fimex_b200
    names = _declared()
    assert len(names) > 30
    missing = [n for n in names if not hasattr(fimex_b200.lib, n)]
    assert not missing, missing
    from fimex_b200 import _lib
    assert not _lib.missing_symbols
    assert set(_lib.SIGNATURES) == set(names)


def test_product_does_not_touch_the_oracle():
    import fimex_b200
    out = subprocess.run(["ldd", fimex_b200.lib_path], capture_output=True, text=True).stdout
    assert "oracle" not in out and "fimexref" not in out
    for dirpath, _, files in os.walk(os.path.join(ROOT, "fimex_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".h", ".cuh", ".cc", ".cpp")):
                src = open(os.path.join(dirpath, f), errors="replace").read()
                assert "oracle_lib" not in src and "liboracle" not in src and "fimex_oracle.h" not in src, f


def test_no_cpu_fallback_without_gpu():
    import fimex_b200
    if fimex_b200.lib.fimexcu_device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(fimex_b200.FimexCudaError):
        fimex_b200.CachedInterpolation("x", "y", 1, np.zeros(4), np.zeros(4), 2, 2, 2, 2)
    with pytest.raises(fimex_b200.FimexCudaError):
        fimex_b200.mifi_points2position([1.0], [0.0, 1.0], 0)


def test_sass_is_sm100():
    import fimex_b200
    r = subprocess.run(["cuobjdump", "-lelf", fimex_b200.lib_path], capture_output=True, text=True)
    if r.returncode != 0:
        pytest.skip("cuobjdump not available")
    assert "sm_100a" in r.stdout
