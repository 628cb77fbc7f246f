"""CPU-side checks of the drop-in boundary: liborbx.so loads, exports every symbol include/*.h declares, and refuses to
compute without a CUDA device (no CPU fallback).  No compute call is made here."""
import ctypes as C
import glob
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    names = []
    for hdr in glob.glob(os.path.join(ROOT, "include", "*.h")):
        src = open(hdr).read()
        src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
        names += re.findall(r"\b(orbx_[a-z0-9_]+)\s*\(", src)
    return sorted(set(names))


def test_library_exports_every_declared_symbol(orbx):
    lib = orbx.lib()
    names = _declared_symbols()
    assert len(names) >= 20
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing
    assert lib.orbx_version().startswith(b"orbx")


def test_keypoint_layout_is_cv_keypoint(orbx):
    assert orbx.KP_DTYPE.itemsize == 28
    assert [orbx.KP_DTYPE.fields[f][1] for f in ("x", "y", "size", "angle", "response", "octave", "class_id")] == [0, 4, 8, 12, 16, 20, 24]


def test_host_hamming_matches_reference_bithack(orbx, oracle):
    rng = np.random.default_rng(0)
    a = rng.integers(0, 256, (200, 32), dtype=np.uint8)
    b = rng.integers(0, 256, (200, 32), dtype=np.uint8)
    for i in range(200):
        assert orbx.DescriptorDistance(a[i], b[i]) == oracle.hamming(a[i], b[i])


def test_bad_parameters_are_rejected(orbx):
    for kw in (dict(nfeatures=0), dict(nlevels=0), dict(nlevels=17), dict(scaleFactor=1.0), dict(minThFAST=0),
               dict(iniThFAST=5, minThFAST=7), dict(blur_taps=3), dict(max_batch=0)):
        with pytest.raises(orbx.OrbxError) as e:
            orbx.ORBextractor(**kw)
        assert e.value.code == orbx.ORBX_ERR_ARG


def test_no_cpu_fallback_without_a_device(orbx):
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        pytest.skip("a CUDA device is present")
    with pytest.raises(orbx.OrbxError) as e:
        orbx.ORBextractor(1000)
    assert e.value.code == orbx.ORBX_ERR_CUDA
    assert b"no CPU fallback" in orbx.lib().orbx_last_error()


def test_product_never_references_the_oracle():
    pkg = os.path.join(ROOT, "3_orb_slam3_selfnote_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cc", ".cpp")):
                src = open(os.path.join(dirpath, f), errors="replace").read()
                assert "import oracle" not in src and "liborb_oracle" not in src and "oracle/" not in src, f


def test_static_knn2_rejects_oversize_and_bad_arguments_before_touching_the_device(orbx):
    """orbx_knn2_device_static (the allocation-free entry point): the size limit and the pointer checks come before any CUDA call"""
    lib = orbx.lib()
    hdr = open(os.path.join(ROOT, "include", "orbx.h")).read()
    maxq = int(re.search(r"#define\s+ORBX_KNN2_STATIC_MAX_QUERIES\s+(\d+)", hdr).group(1))
    assert maxq == 16384
    dummy = C.c_void_p(0x1000)   # never dereferenced: every call below is refused first
    rc = lib.orbx_knn2_device_static(dummy, maxq + 1, dummy, 10, dummy, dummy, None, None)
    assert rc == orbx.ORBX_ERR_CAPACITY and b"ORBX_KNN2_STATIC_MAX_QUERIES" in lib.orbx_last_error()
    assert lib.orbx_knn2_device_static(None, 5, dummy, 10, dummy, dummy, None, None) == orbx.ORBX_ERR_ARG
    assert lib.orbx_knn2_device_static(dummy, 5, None, 10, dummy, dummy, None, None) == orbx.ORBX_ERR_ARG
    assert lib.orbx_knn2_device_static(dummy, -1, dummy, 10, dummy, dummy, None, None) == orbx.ORBX_ERR_ARG
    assert lib.orbx_knn2_device_static(None, 0, None, 0, None, None, None, None) == orbx.ORBX_OK


def test_library_exports_nothing_but_the_c_abi(orbx):
    """liborbx.so's dynamic symbol table = the orbx_* entry points (csrc/orbx.map): no kernel stubs, C++ helpers or CUDA runtime leak out"""
    import subprocess
    out = subprocess.run(["nm", "-D", "--defined-only", orbx.LIB_PATH], capture_output=True, text=True, check=True).stdout
    names = [l.split()[-1] for l in out.splitlines() if l.strip()]
    assert names and all(n.startswith("orbx_") for n in names), [n for n in names if not n.startswith("orbx_")][:5]
    assert set(_declared_symbols()) <= set(names)
