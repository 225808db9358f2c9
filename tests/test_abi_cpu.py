"""The C-ABI library loads and exports every symbol include/dmsodom/*_c.h declares (no GPU needed)."""
import ctypes
import os
import re
import subprocess

import pytest

from conftest import ROOT

HEADERS = [os.path.join(ROOT, "include", "dmsodom", n) for n in ("odom_c.h", "frontend_c.h", "nid_c.h", "splat_c.h", "sequence_c.h", "ring_c.h")]


def declared_symbols():
    names = set()
    for h in HEADERS:
        src = open(h).read()
        src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
        src = re.sub(r"typedef[^;]*\(\s*\*\s*\w+\s*\)[^;]*;", "", src)  # function-pointer typedefs are not entry points
        names |= set(re.findall(r"\b(dms_[a-z0-9_]+)\s*\(", src))
    return sorted(names)


def test_header_symbols_are_exported_and_bound():
    from densemonoslam_b200 import _lib
    names = declared_symbols()
    assert len(names) >= 30
    assert os.path.exists(_lib.LIB_PATH), "build densemonoslam_b200/lib/libdmsodom.so first (__graft_entry__.build())"
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/dmsodom but not exported"
        assert n in _lib.SIGNATURES, f"{n} has no ctypes signature in densemonoslam_b200/_lib.py"
    assert sorted(_lib.SIGNATURES) == names
    assert b"dmsodom" in _lib.load().dms_version()


def test_library_has_sm100a_code_and_no_oracle_dependency():
    from densemonoslam_b200 import _lib
    out = subprocess.run(["ldd", _lib.LIB_PATH], capture_output=True, text=True).stdout
    assert "oracle" not in out
    cuobjdump = "/usr/local/cuda/bin/cuobjdump"
    if os.path.exists(cuobjdump):
        elf = subprocess.run([cuobjdump, "-lelf", _lib.LIB_PATH], capture_output=True, text=True).stdout
        assert "sm_100a" in elf, elf


def test_product_package_never_imports_oracle():
    """The product may mention the oracle in comments, but must not include, import, link or load it."""
    pkg = os.path.join(ROOT, "densemonoslam_b200")
    forbidden = ("oracle/", "import oracle", "from oracle", "liboracle", "cpu_odom", "oracle.oracle", "_ref/")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".sh")):
                txt = open(os.path.join(dp, f)).read()
                for pat in forbidden:
                    assert pat not in txt, (os.path.join(dp, f), pat)


def test_create_fails_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from densemonoslam_b200.odometry import RGBDOdometry, DmsError
    with pytest.raises(DmsError):
        RGBDOdometry(640, 480, 320, 240, 528, 528)


def test_headers_compile_as_c_and_cpp(tmp_path):
    """The four C-ABI headers are plain C (and valid C++): no torch / CUDA types in the signatures."""
    inc = os.path.join(ROOT, "include")
    src = '#include "dmsodom/odom_c.h"\n#include "dmsodom/frontend_c.h"\n#include "dmsodom/splat_c.h"\n#include "dmsodom/nid_c.h"\n#include "dmsodom/sequence_c.h"\nint main(void) { return 0; }\n'
    for comp, name, std in (("gcc", "t.c", "-std=c99"), ("g++", "t.cpp", "-std=c++11")):
        f = tmp_path / name
        f.write_text(src)
        r = subprocess.run([comp, std, "-Wall", "-Werror", "-fsyntax-only", "-I", inc, str(f)], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr


def test_threshold_equivalents_for_every_kind_of_threshold():
    """The squared-domain thresholds must terminate and be exact for ordinary, huge and infinite thresholds (a caller may
    pass +inf to switch a test off)."""
    import ctypes as C
    import numpy as np
    from densemonoslam_b200 import _lib
    L = _lib.load()
    f32 = np.float32
    for T in (0.0, 1e-30, 0.1, 0.3420201, 1.0, 12345.678, 1e19, 3e19, float(np.finfo(f32).max), float("inf")):
        d, a = C.c_float(), C.c_float()
        L.dms_threshold_equiv(T, T, C.byref(d), C.byref(a))
        d, a, T32 = f32(d.value), f32(a.value), f32(T)
        with np.errstate(over="ignore", invalid="ignore"):
            # d = the largest float whose sqrt is <= T; a = the smallest float whose sqrt is >= T
            assert np.sqrt(d) <= T32 and (np.isinf(d) or not np.sqrt(np.nextafter(d, f32(np.inf))) <= T32), (T, d)
            if T > 0:
                assert (np.isinf(a) or np.sqrt(a) >= T32) and (a == 0 or np.sqrt(np.nextafter(a, f32(0))) < T32), (T, a)
    d, a = C.c_float(), C.c_float()
    L.dms_threshold_equiv(float("nan"), -1.0, C.byref(d), C.byref(a))
    assert d.value < 0 and a.value == 0   # nothing passes
