"""CPU suite, part 3: the C-ABI library loads, exports every symbol include/orbx.h declares, and fails loudly
(no CPU fallback) when there is no CUDA device.  No compute calls are made without a GPU."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import orb_slam_cuda_b200 as X

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "orbx.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(orbx_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    names = _declared()
    assert len(names) >= 25 and "orbx_extract" in names and "orbx_stereo_match" in names
    lib = X.lib()
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, f"declared in include/orbx.h but not exported by liborbx.so: {missing}"


def test_keypoint_layout_matches_cv_keypoint():
    assert X.KP_DTYPE.itemsize == 28
    assert [X.KP_DTYPE.fields[f][1] for f in ("x", "y", "size", "angle", "response", "octave", "class_id")] == [0, 4, 8, 12, 16, 20, 24]
    assert C.sizeof(X.Params) == 36


def test_version_and_stage_names():
    lib = X.lib()
    assert lib.orbx_version() >= 100
    lib.orbx_stage_name.restype = C.c_char_p
    names = [lib.orbx_stage_name(i).decode() for i in range(5)]
    # the reference's GetTime rows (src/ORBextractor.cc:1841, :1131, :892, :1737, :1331/:1753)
    assert names[0] == "Pyramid/Resize" and names[1] == "FAST+Grid" and names[2] == "Make quadtree" and names[3] == "Gaussian Blur"


def test_bad_arguments_are_rejected_without_a_device():
    lib = X.lib()
    h = C.c_void_p()
    p = X.Params(0, 1.2, 8, 20, 7, 0, 0, 1, -1)  # nfeatures = 0
    assert lib.orbx_extractor_create(C.byref(p), C.byref(h)) == X.ORBX_E_INVALID
    assert b"nfeatures" in lib.orbx_last_error()
    p = X.Params(1000, 1.0, 8, 20, 7, 0, 0, 1, -1)  # scaleFactor must exceed 1
    assert lib.orbx_extractor_create(C.byref(p), C.byref(h)) == X.ORBX_E_INVALID
    assert lib.orbx_extractor_create(None, C.byref(h)) == X.ORBX_E_INVALID
    n = C.c_int(-1)
    assert lib.orbx_extract(None, None, 0, 0, C.c_size_t(0), None, None, 0, C.byref(n)) != 0  # null handle


def _no_gpu():
    try:
        import torch
        return not torch.cuda.is_available()
    except Exception:
        return True


@pytest.mark.skipif(not _no_gpu(), reason="checks the behaviour of a box without a CUDA device")
def test_no_cpu_fallback():
    with pytest.raises(X.OrbxError) as e:
        X.ORBextractor(1000, 1.2, 8, 20, 7)
    assert e.value.code == X.ORBX_E_CUDA and "no CPU path" in str(e.value)
    with pytest.raises(X.OrbxError) as e:
        X.ORBmatcher.DescriptorDistance(np.zeros(32, np.uint8), np.zeros(32, np.uint8))
    assert e.value.code == X.ORBX_E_CUDA
    with pytest.raises(X.OrbxError):
        X.ORBmatcher().match_bruteforce(np.zeros((4, 32), np.uint8), np.zeros((4, 32), np.uint8))


def test_product_does_not_touch_the_oracle():
    """The product package must not import, link or execute anything under oracle/."""
    pkg = os.path.join(ROOT, "orb_slam_cuda_b200")
    for dirpath, _, files in os.walk(pkg):
        if "build" in dirpath.split(os.sep):
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".inc")):
                txt = open(os.path.join(dirpath, f), errors="replace").read()
                assert "oracle_lib" not in txt and "orb_oracle" not in txt and "liborb_oracle" not in txt, os.path.join(dirpath, f)
    for f in os.listdir(os.path.join(ROOT, "include")):
        txt = open(os.path.join(ROOT, "include", f)).read()
        assert "orb_oracle" not in txt
