"""TEST INFRASTRUCTURE: ctypes access to oracle/_ref/libref_dbow2.so — the reference's own DBoW2 compiled from
/root/reference/Thirdparty/DBoW2 by oracle/ref_dbow2/Makefile.  Only tests/ and tools/gen_dbow2_golden.py use it."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(_HERE, "..", "_ref", "libref_dbow2.so")
REF = "/root/reference/Thirdparty/DBoW2"


def build():
    """Compiles the reference sources where they lie; returns the .so path or None when /root/reference is absent (GPU box)."""
    if os.path.isdir(REF):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return SO if os.path.exists(SO) else None


_lib = None


def lib():
    global _lib
    if _lib is None:
        so = build()
        if so is None:
            return None
        _lib = C.CDLL(so)
        _lib.ref_dbow2_load.restype = C.c_void_p
        _lib.ref_dbow2_load.argtypes = [C.c_char_p]
        _lib.ref_dbow2_free.argtypes = [C.c_void_p]
        _lib.ref_dbow2_size.argtypes = [C.c_void_p]
        _lib.ref_dbow2_size.restype = C.c_uint
    return _lib


def _p(a):
    return C.c_void_p(a.ctypes.data)


class RefVocabulary:
    def __init__(self, path):
        self._h = lib().ref_dbow2_load(str(path).encode())
        if not self._h:
            raise RuntimeError(f"reference DBoW2 could not load {path}")

    def size(self):
        return int(lib().ref_dbow2_size(C.c_void_p(self._h)))

    def transform_features(self, desc, levelsup):
        desc = np.ascontiguousarray(desc, dtype=np.uint8)
        n = desc.shape[0]
        word, node, w = np.empty(n, np.uint32), np.empty(n, np.uint32), np.empty(n, np.float64)
        lib().ref_dbow2_transform_features(C.c_void_p(self._h), _p(desc), n, int(levelsup), _p(word), _p(w), _p(node))
        return word, w, node

    def transform(self, desc, levelsup):
        desc = np.ascontiguousarray(desc, dtype=np.uint8)
        n = desc.shape[0]
        bw, bv = np.empty(max(n, 1), np.uint32), np.empty(max(n, 1), np.float64)
        fn, fc, fi = np.empty(max(n, 1), np.uint32), np.empty(max(n, 1), np.int32), np.empty(max(n, 1), np.uint32)
        m = C.c_int(0)
        k = lib().ref_dbow2_transform(C.c_void_p(self._h), _p(desc), n, int(levelsup), _p(bw), _p(bv), C.byref(m), _p(fn), _p(fc), _p(fi))
        bow = {int(bw[i]): float(bv[i]) for i in range(k)}
        fv, o = {}, 0
        for j in range(m.value):
            fv[int(fn[j])] = [int(x) for x in fi[o:o + fc[j]]]
            o += int(fc[j])
        return bow, fv

    def __del__(self):
        try:
            if self._h:
                lib().ref_dbow2_free(C.c_void_p(self._h))
                self._h = None
        except Exception:
            pass
