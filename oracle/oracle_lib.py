"""ctypes bindings to the CPU oracle (oracle/liborb_oracle.so) — TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this module.
The product package (orb_slam_cuda_b200/) never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liborb_oracle.so")

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                     ("octave", "<i4"), ("class_id", "<i4")])
assert KP_DTYPE.itemsize == 28

_u8p = C.POINTER(C.c_uint8)
_f32p = C.POINTER(C.c_float)
_i32p = C.POINTER(C.c_int)


def build(force: bool = False) -> str:
    srcs = [os.path.join(_HERE, f) for f in ("orb_oracle.cpp", "orb_oracle_capi.cpp", "orb_oracle.hpp")]
    if force or not os.path.exists(_SO) or any(os.path.getmtime(s) > os.path.getmtime(_SO) for s in srcs):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
        L = _lib
        L.orc_fast_atan2.restype = C.c_float
        L.orc_fast_atan2.argtypes = [C.c_float, C.c_float]
        L.orc_ic_angle.restype = C.c_float
        L.orc_extractor_create.restype = C.c_void_p
        L.orc_extractor_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
        L.orc_extractor_destroy.argtypes = [C.c_void_p]
        L.orc_bench_extract.restype = C.c_double
        L.orc_bench_bruteforce.restype = C.c_double
    return _lib


def _p(a, t=C.c_void_p):
    return None if a is None else a.ctypes.data_as(t)


def cv_round(v: float) -> int:
    lib().orc_cv_round_f.argtypes = [C.c_float]
    return lib().orc_cv_round_f(C.c_float(v))


def fast_atan2(y, x) -> np.ndarray:
    y = np.ascontiguousarray(y, dtype=np.float32)
    x = np.ascontiguousarray(x, dtype=np.float32)
    out = np.empty_like(y)
    lib().orc_fast_atan2_array(_p(y), _p(x), _p(out), C.c_int(y.size))
    return out


def sincosf(ang) -> tuple[np.ndarray, np.ndarray]:
    ang = np.ascontiguousarray(ang, dtype=np.float32)
    c, s = np.empty_like(ang), np.empty_like(ang)
    lib().orc_sincosf(_p(ang), _p(c), _p(s), C.c_int(ang.size))
    return c, s


def resize_linear(src: np.ndarray, dw: int, dh: int) -> np.ndarray:
    src = np.ascontiguousarray(src, dtype=np.uint8)
    dst = np.empty((dh, dw), dtype=np.uint8)
    lib().orc_resize_linear_8u(_p(src), src.shape[1], src.shape[0], src.strides[0], _p(dst), dw, dh, dw)
    return dst


def border_reflect101(src: np.ndarray, pad: int) -> np.ndarray:
    src = np.ascontiguousarray(src, dtype=np.uint8)
    h, w = src.shape
    dst = np.empty((h + 2 * pad, w + 2 * pad), dtype=np.uint8)
    lib().orc_border_reflect101(_p(src), w, h, src.strides[0], _p(dst), w + 2 * pad, pad)
    return dst


def fast9(img: np.ndarray, threshold: int, nms: bool = True) -> np.ndarray:
    assert img.dtype == np.uint8 and img.strides[1] == 1
    h, w = img.shape
    out = np.empty(max(1, h * w), dtype=KP_DTYPE)
    n = lib().orc_fast9(_p(img), w, h, img.strides[0], threshold, int(nms), _p(out), out.size)
    return out[:n].copy()


def gaussian7(img: np.ndarray) -> np.ndarray:
    assert img.dtype == np.uint8 and img.strides[1] == 1
    h, w = img.shape
    dst = np.empty((h, w), dtype=np.uint8)
    lib().orc_gaussian7(_p(img), w, h, img.strides[0], _p(dst), w)
    return dst


def ic_angle(plane: np.ndarray, x: int, y: int) -> float:
    assert plane.dtype == np.uint8 and plane.strides[1] == 1
    return float(lib().orc_ic_angle(_p(plane), C.c_int(plane.strides[0]), C.c_int(x), C.c_int(y)))


def descriptor(plane: np.ndarray, x: int, y: int, angle: float) -> np.ndarray:
    assert plane.dtype == np.uint8 and plane.strides[1] == 1
    d = np.empty(32, dtype=np.uint8)
    lib().orc_descriptor(_p(plane), C.c_int(plane.strides[0]), C.c_int(x), C.c_int(y), C.c_float(angle), _p(d))
    return d


PATTERN_REFERENCE, PATTERN_CANONICAL = 0, 1


def set_pattern_variant(v: int) -> None:
    """bit_pattern_31_ used by descriptor() / Extractor.extract: PATTERN_REFERENCE (default) = the reference as it compiles
    (element 96 = VX_FAILURE - 2 = -3, src/ORBextractor.cc:261-262), PATTERN_CANONICAL = the ORB pattern of OpenCV / upstream."""
    lib().orc_set_pattern_variant(int(v))


def pattern_values() -> np.ndarray:
    out = np.empty(1024, np.int32)
    lib().orc_pattern_values(_p(out))
    return out


def descriptor_distance(a: np.ndarray, b: np.ndarray) -> int:
    a = np.ascontiguousarray(a, dtype=np.uint8)
    b = np.ascontiguousarray(b, dtype=np.uint8)
    return int(lib().orc_descriptor_distance(_p(a), _p(b)))


def distribute(keys: np.ndarray, minX: int, maxX: int, minY: int, maxY: int, N: int) -> np.ndarray:
    keys = np.ascontiguousarray(keys, dtype=KP_DTYPE)
    out = np.empty(max(keys.size, 1), dtype=KP_DTYPE)
    n = lib().orc_distribute(_p(keys), keys.size, minX, maxX, minY, maxY, N, _p(out), out.size)
    return out[:n].copy()


class Extractor:
    """CPU oracle of ORBextractor (reference src/ORBextractor.cc, CPU path)."""

    def __init__(self, nfeatures=2000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7):
        self.nlevels = nlevels
        self.nfeatures = nfeatures
        self._h = C.c_void_p(lib().orc_extractor_create(nfeatures, C.c_float(scale_factor), nlevels, ini_th, min_th))

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_extractor_destroy(self._h)
            self._h = None

    def tables(self):
        n = self.nlevels
        sf, inv, s2, inv2 = (np.empty(n, np.float32) for _ in range(4))
        quota, umax = np.empty(n, np.int32), np.empty(16, np.int32)
        lib().orc_extractor_tables(self._h, _p(sf), _p(inv), _p(s2), _p(inv2), _p(quota), _p(umax))
        return dict(scale=sf, inv_scale=inv, sigma2=s2, inv_sigma2=inv2, quota=quota, umax=umax)

    def extract(self, img: np.ndarray):
        assert img.dtype == np.uint8 and img.ndim == 2 and img.strides[1] == 1
        h, w = img.shape
        cap = self.nfeatures + 64 * self.nlevels + 64
        kps = np.empty(cap, dtype=KP_DTYPE)
        desc = np.empty((cap, 32), dtype=np.uint8)
        n = lib().orc_extract(self._h, _p(img), w, h, C.c_int(img.strides[0]), _p(kps), _p(desc), cap)
        assert n <= cap
        return kps[:n].copy(), desc[:n].copy()

    def level_size(self, l):
        w, h = C.c_int(), C.c_int()
        lib().orc_level_size(self._h, l, C.byref(w), C.byref(h))
        return w.value, h.value

    def level_padded(self, l) -> np.ndarray:
        w, h = self.level_size(l)
        out = np.empty((h + 38, w + 38), dtype=np.uint8)
        lib().orc_level_padded(self._h, l, _p(out))
        return out

    def level_blurred(self, l):
        w, h = self.level_size(l)
        out = np.empty((h, w), dtype=np.uint8)
        ok = lib().orc_level_blurred(self._h, l, _p(out))
        return out if ok else None

    def _kps(self, fn, l):
        n = fn(self._h, l, None, 0)
        out = np.empty(max(n, 1), dtype=KP_DTYPE)
        fn(self._h, l, _p(out), n)
        return out[:n].copy()

    def level_candidates(self, l):
        return self._kps(lib().orc_level_candidates, l)

    def level_keypoints(self, l):
        return self._kps(lib().orc_level_keypoints, l)


def features_in_area(keys, bounds, x, y, r, min_level=-1, max_level=-1):
    """Frame::GetFeaturesInArea (src/Frame.cc:326-379) over the grid of AssignFeaturesToGrid (:229-244)."""
    keys = np.ascontiguousarray(keys, dtype=KP_DTYPE)
    out = np.empty(max(keys.size, 1), np.int32)
    b = [C.c_float(v) for v in bounds]
    n = lib().orc_features_in_area(_p(keys), keys.size, *b, C.c_float(x), C.c_float(y), C.c_float(r), int(min_level), int(max_level), _p(out))
    return out[:n].copy()


def search_for_initialization(k1, d1, k2, d2, bounds, prev_xy, window=100, nnratio=0.9, check_ori=True):
    k1 = np.ascontiguousarray(k1, dtype=KP_DTYPE)
    k2 = np.ascontiguousarray(k2, dtype=KP_DTYPE)
    d1 = np.ascontiguousarray(d1, dtype=np.uint8)
    d2 = np.ascontiguousarray(d2, dtype=np.uint8)
    prev = np.ascontiguousarray(prev_xy, dtype=np.float32).copy()
    m12 = np.empty(k1.size, dtype=np.int32)
    minX, maxX, minY, maxY = (C.c_float(v) for v in bounds)
    n = lib().orc_search_for_initialization(_p(k1), k1.size, _p(d1), _p(k2), k2.size, _p(d2), minX, maxX, minY, maxY,
                                            _p(prev), _p(m12), C.c_int(window), C.c_float(nnratio), int(check_ori))
    return n, m12, prev


def search_by_projection_frame(x3Dc, valid1, obs1, mp_desc1, octave1, angle1, k2, d2, u_right2, occupied2, cam, mbf, bounds,
                               scale_factors, th, direction=0, check_ori=True, orb_dist=100, depth_test=True):
    x3Dc = np.ascontiguousarray(x3Dc, dtype=np.float32)
    valid1, obs1, occupied2 = (np.ascontiguousarray(a, dtype=np.uint8) for a in (valid1, obs1, occupied2))
    mp_desc1, d2 = (np.ascontiguousarray(a, dtype=np.uint8) for a in (mp_desc1, d2))
    octave1 = np.ascontiguousarray(octave1, dtype=np.int32)
    angle1 = np.ascontiguousarray(angle1, dtype=np.float32)
    k2 = np.ascontiguousarray(k2, dtype=KP_DTYPE)
    ur = None if u_right2 is None else np.ascontiguousarray(u_right2, dtype=np.float32)
    cam = np.asarray(cam, dtype=np.float32)
    b = np.asarray(bounds, dtype=np.float32)
    sf = np.ascontiguousarray(scale_factors, dtype=np.float32)
    m21 = np.empty(max(k2.size, 1), np.int32)
    n = lib().orc_search_by_projection_frame(valid1.size, _p(x3Dc), _p(valid1), _p(obs1), _p(mp_desc1), _p(octave1), _p(angle1),
                                             _p(k2), k2.size, _p(d2), _p(ur) if ur is not None else None, _p(occupied2),
                                             _p(cam), C.c_float(mbf), _p(b), _p(sf), C.c_float(th), int(direction),
                                             int(check_ori), _p(m21), int(orb_dist), int(depth_test))
    return n, m21[:k2.size]


def search_by_projection_map(proj, view_cos, level1, valid1, obs1, mp_desc1, k2, d2, u_right2, occupied2, bounds, scale_factors,
                             th, nnratio=0.8):
    proj = np.ascontiguousarray(proj, dtype=np.float32)
    view_cos = np.ascontiguousarray(view_cos, dtype=np.float32)
    level1 = np.ascontiguousarray(level1, dtype=np.int32)
    valid1, obs1, occupied2 = (np.ascontiguousarray(a, dtype=np.uint8) for a in (valid1, obs1, occupied2))
    mp_desc1, d2 = (np.ascontiguousarray(a, dtype=np.uint8) for a in (mp_desc1, d2))
    k2 = np.ascontiguousarray(k2, dtype=KP_DTYPE)
    ur = None if u_right2 is None else np.ascontiguousarray(u_right2, dtype=np.float32)
    b = np.asarray(bounds, dtype=np.float32)
    sf = np.ascontiguousarray(scale_factors, dtype=np.float32)
    m21 = np.empty(max(k2.size, 1), np.int32)
    n = lib().orc_search_by_projection_map(valid1.size, _p(proj), _p(view_cos), _p(level1), _p(valid1), _p(obs1), _p(mp_desc1),
                                           _p(k2), k2.size, _p(d2), _p(ur) if ur is not None else None, _p(occupied2), _p(b),
                                           _p(sf), C.c_float(th), C.c_float(nnratio), _p(m21))
    return n, m21[:k2.size]


def search_by_projection_sim3(uvr, level1, valid1, mp_desc1, k2, d2, occupied2, bounds, th_dist=50):
    uvr = np.ascontiguousarray(uvr, dtype=np.float32)
    level1 = np.ascontiguousarray(level1, dtype=np.int32)
    valid1, occupied2 = (np.ascontiguousarray(a, dtype=np.uint8) for a in (valid1, occupied2))
    mp_desc1, d2 = (np.ascontiguousarray(a, dtype=np.uint8) for a in (mp_desc1, d2))
    k2 = np.ascontiguousarray(k2, dtype=KP_DTYPE)
    b = np.asarray(bounds, dtype=np.float32)
    m21 = np.empty(max(k2.size, 1), np.int32)
    n = lib().orc_search_by_projection_sim3(valid1.size, _p(uvr), _p(level1), _p(valid1), _p(mp_desc1), _p(k2), k2.size, _p(d2),
                                            _p(occupied2), _p(b), int(th_dist), _p(m21))
    return n, m21[:k2.size]


def _bow_args(start1, idx1, start2, idx2, valid1, d1, angle1, valid2, d2, angle2):
    start1, start2 = (np.ascontiguousarray(a, dtype=np.int32) for a in (start1, start2))
    idx1, idx2 = (np.ascontiguousarray(a, dtype=np.uint32) for a in (idx1, idx2))
    valid1, valid2 = (np.ascontiguousarray(a, dtype=np.uint8) for a in (valid1, valid2))
    d1, d2 = (np.ascontiguousarray(a, dtype=np.uint8) for a in (d1, d2))
    angle1, angle2 = (np.ascontiguousarray(a, dtype=np.float32) for a in (angle1, angle2))
    return start1, idx1, start2, idx2, valid1, d1, angle1, valid2, d2, angle2


def search_by_bow(start1, idx1, start2, idx2, valid1, d1, angle1, valid2, d2, angle2, th_dist=50, nnratio=0.7, check_ori=True):
    start1, idx1, start2, idx2, valid1, d1, angle1, valid2, d2, angle2 = _bow_args(start1, idx1, start2, idx2, valid1, d1, angle1,
                                                                                   valid2, d2, angle2)
    n1, n2 = valid1.size, valid2.size
    m12, m21 = np.empty(max(n1, 1), np.int32), np.empty(max(n2, 1), np.int32)
    n = lib().orc_search_by_bow(start1.size - 1, _p(start1), _p(idx1), _p(start2), _p(idx2), n1, _p(valid1), _p(d1), _p(angle1), n2,
                                _p(valid2), _p(d2), _p(angle2), int(th_dist), C.c_float(nnratio), int(check_ori), _p(m12), _p(m21))
    return n, m12[:n1], m21[:n2]


def fuse_match(uvr, ur1, level1, valid1, mp_desc1, k2, d2, u_right2, inv_level_sigma2, bounds):
    uvr = np.ascontiguousarray(uvr, dtype=np.float32)
    ur1 = None if ur1 is None else np.ascontiguousarray(ur1, dtype=np.float32)
    level1 = np.ascontiguousarray(level1, dtype=np.int32)
    valid1, mp_desc1, d2 = (np.ascontiguousarray(a, dtype=np.uint8) for a in (valid1, mp_desc1, d2))
    k2 = np.ascontiguousarray(k2, dtype=KP_DTYPE)
    u_right2 = None if u_right2 is None else np.ascontiguousarray(u_right2, dtype=np.float32)
    sg = np.ascontiguousarray(inv_level_sigma2, dtype=np.float32)
    b = np.asarray(bounds, dtype=np.float32)
    n1 = valid1.size
    idx, dist = np.empty(max(n1, 1), np.int32), np.empty(max(n1, 1), np.int32)
    lib().orc_fuse_match(n1, _p(uvr), _p(ur1), _p(level1), _p(valid1), _p(mp_desc1), _p(k2), k2.size, _p(d2), _p(u_right2), _p(sg), _p(b),
                         _p(idx), _p(dist))
    return idx[:n1], dist[:n1]


def search_for_triangulation(start1, idx1, start2, idx2, k1, d1, valid1, stereo1, k2, d2, valid2, stereo2, F12, ex, ey, scale_factors2,
                             level_sigma2_2, check_ori=True):
    start1, start2 = (np.ascontiguousarray(a, dtype=np.int32) for a in (start1, start2))
    idx1, idx2 = (np.ascontiguousarray(a, dtype=np.uint32) for a in (idx1, idx2))
    k1, k2 = (np.ascontiguousarray(a, dtype=KP_DTYPE) for a in (k1, k2))
    d1, d2, valid1, valid2, stereo1, stereo2 = (np.ascontiguousarray(a, dtype=np.uint8) for a in (d1, d2, valid1, valid2, stereo1, stereo2))
    F = np.ascontiguousarray(F12, dtype=np.float32).reshape(9)
    sf, sg = (np.ascontiguousarray(a, dtype=np.float32) for a in (scale_factors2, level_sigma2_2))
    m12 = np.empty(max(k1.size, 1), np.int32)
    n = lib().orc_search_for_triangulation(max(start1.size - 1, 0), _p(start1), _p(idx1), _p(start2), _p(idx2), k1.size, _p(k1), _p(d1),
                                           _p(valid1), _p(stereo1), k2.size, _p(k2), _p(d2), _p(valid2), _p(stereo2), _p(F),
                                           C.c_float(ex), C.c_float(ey), _p(sf), _p(sg), int(check_ori), _p(m12))
    return n, m12[:k1.size]


def cvt_color_gray(img, bgr=False, cv2_coeffs=False):
    """cv::cvtColor 8-bit RGB(A)/BGR(A) -> gray in OpenCV's fixed point (numpy restatement; pinned against cv2 4.13 for the default
    coefficients in tests/test_color.py; the 14-bit set is OpenCV 2.4 / 3.x's published constants)."""
    img = np.asarray(img, dtype=np.uint8)
    c = [img[..., i].astype(np.int64) for i in range(3)]
    r, g, b = (c[2], c[1], c[0]) if bgr else (c[0], c[1], c[2])
    if cv2_coeffs:
        return ((r * 4899 + g * 9617 + b * 1868 + (1 << 13)) >> 14).astype(np.uint8)
    return ((r * 9798 + g * 19235 + b * 3735 + (1 << 14)) >> 15).astype(np.uint8)


def undistort_keypoints(kps, K, dist):
    """Frame::UndistortKeyPoints (src/Frame.cc:403-433) = cv::undistortPoints(mat, mat, mK, mDistCoef, cv::Mat(), mK): OpenCV's
    cvUndistortPointsInternal restated in numpy float64 (individually rounded operations in its order: normalise with 1/fx, five
    iterations of the inverse distortion model, re-project with P = K, one rounding to float).  Pinned against cv2.undistortPoints
    (tests/test_undistort.py).  dist[0] == 0: the reference's early return, keypoints unchanged."""
    kps = np.ascontiguousarray(kps, dtype=KP_DTYPE).copy()
    d = np.zeros(5, np.float64)
    dd = np.asarray(dist, np.float32).reshape(-1)
    d[:dd.size] = dd.astype(np.float64)
    if dd.size == 0 or dd[0] == 0:
        return kps
    fx, fy, cx, cy = (np.float64(np.float32(v)) for v in K)
    k1, k2, p1, p2, k3 = d
    ifx, ify = 1.0 / fx, 1.0 / fy
    u, v = kps["x"].astype(np.float64), kps["y"].astype(np.float64)
    x, y = (u - cx) * ifx, (v - cy) * ify
    x0, y0 = x.copy(), y.copy()
    frozen = np.zeros(x.shape, bool)
    for _ in range(5):
        r2 = x * x + y * y
        icdist = 1.0 / (1.0 + ((k3 * r2 + k2) * r2 + k1) * r2)
        neg = (icdist < 0) & ~frozen
        deltaX = 2 * p1 * x * y + p2 * (r2 + 2 * x * x)
        deltaY = p1 * (r2 + 2 * y * y) + 2 * p2 * x * y
        nx, ny = (x0 - deltaX) * icdist, (y0 - deltaY) * icdist
        nx = np.where(neg, (u - cx) * ifx, nx)
        ny = np.where(neg, (v - cy) * ify, ny)
        x, y = np.where(frozen, x, nx), np.where(frozen, y, ny)
        frozen |= neg
    xx, yy, ww = fx * x + 0.0 * y + cx, 0.0 * x + fy * y + cy, 1.0 / (0.0 * x + 0.0 * y + 1.0)
    kps["x"], kps["y"] = (xx * ww).astype(np.float32), (yy * ww).astype(np.float32)
    return kps


def distinctive_descriptors(start, desc):
    start = np.ascontiguousarray(start, dtype=np.int32)
    desc = np.ascontiguousarray(desc, dtype=np.uint8)
    n = max(start.size - 1, 0)
    idx, med = np.empty(max(n, 1), np.int32), np.empty(max(n, 1), np.int32)
    lib().orc_distinctive_descriptors(n, _p(start), _p(desc), _p(idx), _p(med))
    return idx[:n], med[:n]


class Vocabulary:
    """The node table DBoW2's loadFromTextFile builds (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1338-1424): node 0 = root,
    then one node per line `parent is_leaf d0 .. d31 weight`; first line `k L scoring weighting`."""

    def __init__(self, k, L, scoring, weighting, parent, is_leaf, desc, weight):
        self.k, self.L, self.scoring, self.weighting = int(k), int(L), int(scoring), int(weighting)
        self.parent = np.ascontiguousarray(parent, dtype=np.int32)
        self.is_leaf = np.ascontiguousarray(is_leaf, dtype=np.uint8)
        self.desc = np.ascontiguousarray(desc, dtype=np.uint8)
        self.weight = np.ascontiguousarray(weight, dtype=np.float64)

    @classmethod
    def from_text(cls, path):
        """loadFromTextFile, :1338-1424: one node per line.  The reference's `while(!f.eof()) getline` also runs on the empty
        remainder after the last newline, where `ssnode >> pid` fails in the sentry and pid / nIsLeaf stay uninitialised
        (undefined behaviour; checked against the compiled reference: stale value or crash) — blank lines make no node here."""
        with open(path) as f:
            text = f.read()
        lines = text.split("\n")
        k, L, n1, n2 = (int(t) for t in lines[0].split())
        parent, leaf, desc, weight = [-1], [0], [[0] * 32], [0.0]
        for line in lines[1:]:
            t = line.split()
            if not t:
                continue
            parent.append(int(t[0]))
            leaf.append(int(t[1]) > 0)
            desc.append([int(x) & 0xFF for x in t[2:34]])
            weight.append(float(t[34]))
        return cls(k, L, n1, n2, parent, leaf, desc, weight)

    def to_text(self, path, trailing_newline=True):
        """saveToTextFile's format (:1429-1450); it ends every line, the last one too, with a newline."""
        rows = [f"{self.k} {self.L} {self.scoring} {self.weighting}"]
        for i in range(1, self.parent.size):
            rows.append(f"{self.parent[i]} {int(self.is_leaf[i])} " + " ".join(str(int(b)) for b in self.desc[i]) + f" {float(self.weight[i])!r}")
        with open(path, "w") as f:
            f.write("\n".join(rows) + ("\n" if trailing_newline else ""))

    def transform_features(self, feat, levelsup=4):
        """transform(feature, word_id, weight, &nid, levelsup), :1218-1261, for every row of feat."""
        feat = np.ascontiguousarray(feat, dtype=np.uint8)
        n = feat.shape[0]
        word, node, w = np.empty(max(n, 1), np.uint32), np.empty(max(n, 1), np.uint32), np.empty(max(n, 1), np.float64)
        lib().orc_vocabulary_transform(self.parent.size, _p(self.parent), _p(self.is_leaf), _p(self.desc), _p(self.weight), self.L, n,
                                       _p(feat), int(levelsup), _p(word), _p(w), _p(node))
        return word[:n], w[:n], node[:n]

    def transform(self, feat, levelsup=4):
        """transform(features, BowVector&, FeatureVector&, levelsup), :1127-1194, with BowVector::addWeight / addIfNotExist /
        normalize (Thirdparty/DBoW2/DBoW2/BowVector.cpp:33-88).  Returns (bow: {word: value}, fv: {node: [feature indices]}).
        A descent that ends above the node level leaves `nid` unwritten in the reference; the value it then adds is the one of
        the previous feature (the variable's stack slot), 0 before any — restated as such."""
        word, w, node = self.transform_features(feat, levelsup)
        bow, fv, nid = {}, {}, 0
        tf = self.weighting in (0, 1)                    # TF_IDF, TF
        for i in range(word.size):
            if node[i] != 0xFFFFFFFF:
                nid = int(node[i])
            if w[i] > 0:
                if tf:
                    bow[int(word[i])] = bow.get(int(word[i]), 0.0) + float(w[i])
                else:
                    bow.setdefault(int(word[i]), float(w[i]))
                fv.setdefault(nid, []).append(i)
        must, l1 = self.scoring != 5, self.scoring != 1   # all but DOT_PRODUCT normalise, L2 only for L2_NORM (ScoringObject.h:74-89)
        if tf and bow and not must:
            nd = float(len(bow))
            bow = {k: v / nd for k, v in bow.items()}
        if must:
            keys = sorted(bow)
            norm = 0.0
            for k in keys:
                norm += abs(bow[k]) if l1 else bow[k] * bow[k]
            if not l1:
                norm = float(np.sqrt(norm))
            if norm > 0.0:
                bow = {k: bow[k] / norm for k in keys}
        return dict(sorted(bow.items())), dict(sorted(fv.items()))


def match_bruteforce(q, d, th=50, ratio=0.6):
    q = np.ascontiguousarray(q, dtype=np.uint8)
    d = np.ascontiguousarray(d, dtype=np.uint8)
    nq, nd = q.shape[0], d.shape[0]
    idx, d1, d2, m = (np.empty(nq, np.int32) for _ in range(4))
    lib().orc_match_bruteforce(_p(q), nq, _p(d), nd, th, C.c_float(ratio), _p(idx), _p(d1), _p(d2), _p(m))
    return idx, d1, d2, m


def stereo_matches(exL: Extractor, exR: Extractor, kL, dL, kR, dR, bf: float, fx: float):
    kL = np.ascontiguousarray(kL, dtype=KP_DTYPE)
    kR = np.ascontiguousarray(kR, dtype=KP_DTYPE)
    dL = np.ascontiguousarray(dL, dtype=np.uint8)
    dR = np.ascontiguousarray(dR, dtype=np.uint8)
    u = np.empty(kL.size, np.float32)
    z = np.empty(kL.size, np.float32)
    lib().orc_stereo_matches(exL._h, exR._h, _p(kL), kL.size, _p(dL), _p(kR), kR.size, _p(dR), C.c_float(bf),
                             C.c_float(fx), _p(u), _p(z))
    return u, z


def bench_extract(frames: np.ndarray, nfeatures, scale_factor, nlevels, ini_th, min_th, threads: int):
    frames = np.ascontiguousarray(frames, dtype=np.uint8)
    n, h, w = frames.shape
    nk = C.c_longlong(0)
    lib().orc_bench_extract.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int, C.c_int,
                                        C.c_int, C.c_int, C.POINTER(C.c_longlong)]
    secs = lib().orc_bench_extract(_p(frames), n, w, h, nfeatures, C.c_float(scale_factor), nlevels, ini_th, min_th,
                                   threads, C.byref(nk))
    return secs, nk.value


def bench_bruteforce(q, d, threads: int):
    q = np.ascontiguousarray(q, dtype=np.uint8)
    d = np.ascontiguousarray(d, dtype=np.uint8)
    nq, nd = q.shape[0], d.shape[0]
    idx, d1, d2, m = (np.empty(nq, np.int32) for _ in range(4))
    secs = lib().orc_bench_bruteforce(_p(q), nq, _p(d), nd, threads, _p(idx), _p(d1), _p(d2), _p(m))
    return secs, (idx, d1, d2, m)
