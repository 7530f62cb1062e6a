"""TEST INFRASTRUCTURE: ctypes access to oracle/_ref/libref_extractor.so — the reference's own src/ORBextractor.cc compiled from
/root/reference by oracle/ref_extractor/Makefile (CPU branch of operator(); do-nothing OpenVX; the five OpenCV pixel primitives
are the oracle's cv2-pinned models).  Only tests/ and tools/gen_extractor_golden.py use it."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(_HERE, "..", "_ref", "libref_extractor.so")
REF = "/root/reference/src/ORBextractor.cc"
KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"), ("octave", "<i4"),
                     ("class_id", "<i4")])


def build():
    """Compiles the reference source where it lies; returns the .so path or None when /root/reference is absent (GPU box)."""
    if os.path.exists(REF):
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
        _lib.refx_create.restype = C.c_void_p
    return _lib


def available():
    return lib() is not None


def _p(a):
    return C.c_void_p(a.ctypes.data)


class RefExtractor:
    """ORB_SLAM2::ORBextractor of the reference (src/ORBextractor.cc:496-560 constructor, :1538-1815 operator())."""

    def __init__(self, width, height, nfeatures=2000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7):
        self.nlevels, self.nfeatures = nlevels, nfeatures
        self._h = C.c_void_p(lib().refx_create(nfeatures, C.c_float(scale_factor), nlevels, ini_th, min_th, width, height))

    def __del__(self):
        if getattr(self, "_h", None):
            lib().refx_destroy(self._h)
            self._h = None

    def tables(self):
        quota, scale = np.zeros(self.nlevels, np.int32), np.zeros(self.nlevels, np.float32)
        umax, pat = np.zeros(16, np.int32), np.zeros(1024, np.int32)
        lib().refx_tables(self._h, _p(quota), _p(scale), _p(umax), _p(pat))
        return dict(quota=quota, scale=scale, umax=umax, pattern=pat)

    def extract(self, img):
        assert img.dtype == np.uint8 and img.ndim == 2 and img.strides[1] == 1
        h, w = img.shape
        cap = self.nfeatures + 64 * self.nlevels + 64
        k7 = np.empty((cap, 7), np.float32)
        desc = np.empty((cap, 32), np.uint8)
        n = lib().refx_extract(self._h, _p(img), w, h, C.c_size_t(img.strides[0]), _p(k7), _p(desc), cap)
        assert 0 <= n <= cap, n
        kps = np.zeros(n, KP_DTYPE)
        for i, f in enumerate(("x", "y", "size", "angle", "response")):
            kps[f] = k7[:n, i]
        kps["octave"] = k7[:n, 5].astype(np.int32)
        kps["class_id"] = k7[:n, 6].astype(np.int32)
        return kps, desc[:n].copy()

    def level(self, l):
        cap = 1 << 24
        out = np.empty(cap, np.uint8)
        r = lib().refx_level(self._h, l, _p(out), cap)
        assert r >= 0
        w, h = r & 0xFFFF, r >> 16
        return out[: w * h].reshape(h, w).copy()

    def distribute(self, keys, minX, maxX, minY, maxY, N, level=0):
        keys = np.ascontiguousarray(keys, dtype=KP_DTYPE)
        xyr = np.stack([keys["x"], keys["y"], keys["response"]], axis=1).astype(np.float32).copy()
        out = np.empty((max(len(keys), 1), 3), np.float32)
        n = lib().refx_distribute(self._h, _p(xyr), len(keys), minX, maxX, minY, maxY, N, level, _p(out), len(out))
        assert n >= 0
        return out[:n].copy()


def bench_extract(frames, nfeatures, scale_factor, nlevels, ini_th, min_th, threads):
    """bench.py's CPU arm: wall seconds for `frames` (n, h, w) on `threads` host threads, one reference extractor per thread."""
    frames = np.ascontiguousarray(frames, dtype=np.uint8)
    n, h, w = frames.shape
    nk = C.c_longlong(0)
    f = lib().refx_bench_extract
    f.restype = C.c_double
    f.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_longlong)]
    secs = f(_p(frames), n, w, h, nfeatures, C.c_float(scale_factor), nlevels, ini_th, min_th, threads, C.byref(nk))
    return secs, nk.value
