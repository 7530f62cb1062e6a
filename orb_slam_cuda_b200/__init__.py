"""orb_slam_cuda_b200 — B200-native ORB-SLAM2 front end (ORBextractor + Hamming matchers + stereo matching).

The product is `liborbx.so` (hand-written sm_100a CUDA kernels behind the C ABI of include/orbx.h) plus the C++
classes of include/ORBextractor.h / ORBmatcher.h that keep the reference's class surface.  This Python package is a
thin ctypes binding of the same C ABI, used by the tests and bench.py; it mirrors the reference names
(ORBextractor.__call__ = operator(), GetScaleFactors, ..., ORBmatcher.DescriptorDistance, SearchForInitialization).

There is no CPU fallback: importing works anywhere, but every compute call fails with OrbxError when the shared
library is missing or no B200 is present.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("ORBX_LIB") or os.path.join(_HERE, "liborbx.so")  # ORBX_LIB: kernel-variant experiments

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                     ("octave", "<i4"), ("class_id", "<i4")])
assert KP_DTYPE.itemsize == 28

ORBX_OK, ORBX_E_INVALID, ORBX_E_CUDA, ORBX_E_OOM, ORBX_E_OVERFLOW = 0, -1, -2, -3, -4
TH_LOW, TH_HIGH, HISTO_LENGTH = 50, 100, 30  # ORBmatcher constants, reference src/ORBmatcher.cc:37-39


class OrbxError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"orbx error {code}: {msg}")
        self.code = code


class Params(C.Structure):
    _fields_ = [("nfeatures", C.c_int), ("scale_factor", C.c_float), ("nlevels", C.c_int), ("ini_th_fast", C.c_int),
                ("min_th_fast", C.c_int), ("width", C.c_int), ("height", C.c_int), ("max_batch", C.c_int),
                ("device", C.c_int)]


_lib = None


def lib() -> C.CDLL:
    """Loads liborbx.so (built in-tree by orb_slam_cuda_b200/build.py).  Fails loudly when it is missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise OrbxError(ORBX_E_CUDA, f"{LIB_PATH} is missing: run `python orb_slam_cuda_b200/build.py` "
                                         "(there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        L.orbx_last_error.restype = C.c_char_p
        L.orbx_kernel_launches.restype = C.c_longlong
        L.orbx_extractor_stream.restype = C.c_void_p
        L.orbx_extractor_stream.argtypes = [C.c_void_p]
        L.orbx_extractor_destroy.argtypes = [C.c_void_p]
        L.orbx_extractor_destroy.restype = None
        L.orbx_extract_batch_wait.argtypes = [C.c_void_p]
        L.orbx_vocabulary_destroy.argtypes = [C.c_void_p]
        L.orbx_vocabulary_destroy.restype = None
        L.orbx_vocabulary_info.argtypes = [C.c_void_p] + [C.c_void_p] * 6
        L.orbx_vocabulary_transform.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        _lib = L
    return _lib


def _check(rc: int) -> None:
    if rc != 0:
        raise OrbxError(rc, lib().orbx_last_error().decode(errors="replace"))


def _p(a):
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return C.c_void_p(a.ctypes.data)
    return C.c_void_p(int(a))  # raw address (e.g. torch tensor .data_ptr())


def kernel_launches() -> int:
    return int(lib().orbx_kernel_launches())


class ORBextractor:
    """Mirror of ORB_SLAM2::ORBextractor (reference include/ORBextractor.h:76-160).

    ORBextractor(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST[, width, height]);
    keypoints, descriptors = extractor(image)      # operator()(image, mask, keypoints, descriptors)
    """

    HARRIS_SCORE, FAST_SCORE = 0, 1

    def __init__(self, nfeatures: int, scaleFactor: float, nlevels: int, iniThFAST: int, minThFAST: int,
                 width: int = 0, height: int = 0, max_batch: int = 1, device: int = -1):
        self._h = C.c_void_p()
        self.nfeatures, self.nlevels, self.scaleFactor = nfeatures, nlevels, float(np.float32(scaleFactor))
        p = Params(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, width, height, max_batch, device)
        _check(lib().orbx_extractor_create(C.byref(p), C.byref(self._h)))
        n = nlevels
        self._sf, self._inv, self._s2, self._inv2 = (np.empty(n, np.float32) for _ in range(4))
        self._quota = np.empty(n, np.int32)
        _check(lib().orbx_get_scale_tables(self._h, _p(self._sf), _p(self._inv), _p(self._s2), _p(self._inv2), _p(self._quota)))

    def __del__(self):
        h = getattr(self, "_h", None)
        if h and _lib is not None:
            _lib.orbx_extractor_destroy(h)
            self._h = None

    # ---- accessors (include/ORBextractor.h:94-114) ----
    def GetLevels(self): return self.nlevels
    def GetScaleFactor(self): return self.scaleFactor
    def GetScaleFactors(self): return self._sf.copy()
    def GetInverseScaleFactors(self): return self._inv.copy()
    def GetScaleSigmaSquares(self): return self._s2.copy()
    def GetInverseScaleSigmaSquares(self): return self._inv2.copy()
    def features_per_level(self): return self._quota.copy()

    PATTERN_REFERENCE, PATTERN_CANONICAL = 0, 1

    def set_pattern(self, variant: int):
        """orbx_extractor_set_pattern: PATTERN_REFERENCE (default) = bit_pattern_31_ as the reference compiles it (element 96 = VX_FAILURE - 2
        = -3, src/ORBextractor.cc:261-262), PATTERN_CANONICAL = the canonical ORB table."""
        lib().orbx_extractor_set_pattern.argtypes = [C.c_void_p, C.c_int]
        _check(lib().orbx_extractor_set_pattern(self._h, int(variant)))

    def umax(self):
        u = np.empty(16, np.int32)
        _check(lib().orbx_get_umax(self._h, _p(u)))
        return u

    @property
    def stream(self) -> int:
        return int(lib().orbx_extractor_stream(self._h) or 0)

    def default_cap(self) -> int:
        return self.nfeatures + 4 * self.nlevels

    # ---- per-stage timing (the reference's GetTime rows) ----
    def enable_timing(self, on: bool = True):
        _check(lib().orbx_extractor_enable_timing(self._h, int(on)))

    def stage_ms(self) -> dict:
        ms = np.zeros(5, np.float32)
        _check(lib().orbx_extractor_stage_ms(self._h, _p(ms)))
        lib().orbx_stage_name.restype = C.c_char_p
        return {lib().orbx_stage_name(i).decode(): float(ms[i]) for i in range(5)}

    # ---- operator() ----
    def __call__(self, image: np.ndarray, mask=None):
        """image: uint8 [h, w] (mask is ignored, like the reference) -> (keypoints KP_DTYPE[n], descriptors uint8[n, 32])"""
        if image is None or image.size == 0:
            return np.empty(0, KP_DTYPE), np.empty((0, 32), np.uint8)
        assert image.dtype == np.uint8 and image.ndim == 2 and image.strides[1] == 1, "CV_8UC1 image expected"
        h, w = image.shape
        sc = getattr(self, "_single", None)
        if sc is None:  # per-extractor scratch for the per-frame call: output arrays, their addresses, the bound C function
            cap = self.default_cap()
            kps, desc, n = np.empty(cap, KP_DTYPE), np.empty((cap, 32), np.uint8), C.c_int(0)
            fn = lib().orbx_extract
            fn.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
            sc = self._single = (fn, kps, desc, n, kps.ctypes.data, desc.ctypes.data, C.addressof(n), cap)
        fn, kps, desc, n, pk, pd, pn, cap = sc
        rc = fn(self._h, image.ctypes.data, w, h, image.strides[0], pk, pd, cap, pn)
        if rc:
            _check(rc)
        m = n.value
        return kps[:m].copy(), desc[:m].copy()

    def extract_batch(self, frames: np.ndarray, kps=None, desc=None, counts=None, cap=None):
        """frames: uint8 [B, h, w] host array (pinned or pageable) -> (kps [B, cap], desc [B, cap, 32], counts [B])"""
        assert frames.dtype == np.uint8 and frames.ndim == 3 and frames.strides[2] == 1
        b, h, w = frames.shape
        cap = cap or self.default_cap()
        kps = np.empty((b, cap), KP_DTYPE) if kps is None else kps
        desc = np.empty((b, cap, 32), np.uint8) if desc is None else desc
        counts = np.empty(b, np.int32) if counts is None else counts
        _check(lib().orbx_extract_batch(self._h, _p(frames), b, w, h, C.c_size_t(frames.strides[1]), C.c_size_t(frames.strides[0]),
                                        _p(kps), _p(desc), cap, _p(counts)))
        return kps, desc, counts

    def extract_batch_async(self, frames: np.ndarray, kps, desc, counts, cap=None):
        """orbx_extract_batch_async: enqueue one batch (all arrays pinned host memory, frames tightly packed) and return; at
        most two calls in flight, each with its own output arrays.  extract_batch_wait() collects the oldest one."""
        assert frames.dtype == np.uint8 and frames.ndim == 3 and frames.strides[2] == 1
        b, h, w = frames.shape
        cap = cap or self.default_cap()
        _check(lib().orbx_extract_batch_async(self._h, _p(frames), b, w, h, C.c_size_t(frames.strides[1]), C.c_size_t(frames.strides[0]),
                                              _p(kps), _p(desc), cap, _p(counts)))

    def extract_batch_wait(self):
        _check(lib().orbx_extract_batch_wait(self._h))

    def extract_batch_device(self, d_frames: int, batch: int, w: int, h: int, d_kps: int, d_desc: int, cap: int,
                             d_counts: int, d_status: int = 0, stream: int = 0, stride: int = 0, frame_stride: int = 0):
        """All arguments are raw device addresses; asynchronous on `stream` (0 = the extractor's own stream)."""
        stride = stride or w
        frame_stride = frame_stride or stride * h
        _check(lib().orbx_extract_batch_device(self._h, _p(d_frames), batch, w, h, C.c_size_t(stride), C.c_size_t(frame_stride),
                                               _p(d_kps), _p(d_desc), cap, _p(d_counts), _p(d_status) if d_status else None,
                                               _p(stream) if stream else None))

    # ---- mvImagePyramid (include/ORBextractor.h:116) ----
    def level_size(self, level: int):
        w, h = C.c_int(), C.c_int()
        _check(lib().orbx_level_size(self._h, level, C.byref(w), C.byref(h)))
        return w.value, h.value

    def image_pyramid_level(self, level: int, frame: int = 0, padded: bool = False) -> np.ndarray:
        """Level `level` of the last extracted frame; padded=True returns the (h+38) x (w+38) parent."""
        ptr, w, h, pitch = C.c_void_p(), C.c_int(), C.c_int(), C.c_size_t()
        _check(lib().orbx_pyramid_level_host(self._h, frame, level, C.byref(ptr), C.byref(w), C.byref(h), C.byref(pitch)))
        e = 19
        base = ptr.value - e * pitch.value - e
        buf = (C.c_uint8 * (pitch.value * (h.value + 2 * e))).from_address(base)
        full = np.frombuffer(buf, np.uint8).reshape(h.value + 2 * e, pitch.value)[:, :w.value + 2 * e]
        return full.copy() if padded else full[e:e + h.value, e:e + w.value].copy()

    # ---- stage taps (parity tests) ----
    def debug_blurred_level(self, level: int, frame: int = 0) -> np.ndarray:
        w, h = self.level_size(level)
        out = np.empty((h, w), np.uint8)
        _check(lib().orbx_debug_blurred_level(self._h, frame, level, _p(out)))
        return out

    def _debug_kps(self, fn, level, frame):
        n = C.c_int(0)
        _check(fn(self._h, frame, level, None, 0, C.byref(n)))
        out = np.empty(max(n.value, 1), KP_DTYPE)
        _check(fn(self._h, frame, level, _p(out), out.size, C.byref(n)))
        return out[:n.value].copy()

    def debug_candidates(self, level: int, frame: int = 0) -> np.ndarray:
        return self._debug_kps(lib().orbx_debug_candidates, level, frame)

    def debug_level_keypoints(self, level: int, frame: int = 0) -> np.ndarray:
        return self._debug_kps(lib().orbx_debug_level_keypoints, level, frame)


def distribute_octtree(keys: np.ndarray, minX: int, maxX: int, minY: int, maxY: int, N: int) -> np.ndarray:
    """Stand-alone DistributeOctTree (reference src/ORBextractor.cc:889-1120) on the device."""
    keys = np.ascontiguousarray(keys, dtype=KP_DTYPE)
    out = np.empty(N + 8, KP_DTYPE)
    n = C.c_int(0)
    _check(lib().orbx_debug_distribute(_p(keys), keys.size, minX, maxX, minY, maxY, N, _p(out), out.size, C.byref(n)))
    return out[:n.value].copy()


class ORBmatcher:
    """Mirror of ORB_SLAM2::ORBmatcher (reference include/ORBmatcher.h:37-102) for the routines on the hot path.

    ORBmatcher(nnratio=0.6, checkOri=True); the static DescriptorDistance, SearchForInitialization and the
    brute-force best / second-best matcher all run on the device through the C ABI (no CPU path)."""

    TH_LOW, TH_HIGH, HISTO_LENGTH = TH_LOW, TH_HIGH, HISTO_LENGTH

    def __init__(self, nnratio: float = 0.6, checkOri: bool = True):
        self.mfNNratio, self.mbCheckOrientation = float(nnratio), bool(checkOri)

    @staticmethod
    def DescriptorDistance(a: np.ndarray, b: np.ndarray):
        """a, b: uint8 [32] -> int, or uint8 [n, 32] -> int32 [n] (row-wise distances); src/ORBmatcher.cc:1647-1663"""
        a = np.ascontiguousarray(a, dtype=np.uint8)
        b = np.ascontiguousarray(b, dtype=np.uint8)
        assert a.shape == b.shape and a.shape[-1] == 32
        n = a.size // 32
        out = np.empty(max(n, 1), np.int32)
        _check(lib().orbx_hamming_distance(_p(a), _p(b), n, _p(out)))
        return int(out[0]) if a.ndim == 1 else out[:n]

    def SearchForInitialization(self, kps1, desc1, kps2, desc2, bounds, vbPrevMatched, windowSize: int = 10):
        """src/ORBmatcher.cc:405-520.  bounds = (mnMinX, mnMaxX, mnMinY, mnMaxY) of the frames; vbPrevMatched is a
        float32 [n1, 2] array updated in place.  Returns (nmatches, vnMatches12 int32 [n1])."""
        kps1 = np.ascontiguousarray(kps1, dtype=KP_DTYPE)
        kps2 = np.ascontiguousarray(kps2, dtype=KP_DTYPE)
        desc1 = np.ascontiguousarray(desc1, dtype=np.uint8)
        desc2 = np.ascontiguousarray(desc2, dtype=np.uint8)
        assert vbPrevMatched.dtype == np.float32 and vbPrevMatched.flags.c_contiguous and vbPrevMatched.size == 2 * kps1.size
        b = np.asarray(bounds, dtype=np.float32)
        m12 = np.empty(max(kps1.size, 1), np.int32)
        n = C.c_int(0)
        _check(lib().orbx_search_for_initialization(_p(kps1), kps1.size, _p(desc1), _p(kps2), kps2.size, _p(desc2), _p(b),
                                                    _p(vbPrevMatched), _p(m12), int(windowSize), C.c_float(self.mfNNratio),
                                                    int(self.mbCheckOrientation), C.byref(n)))
        return n.value, m12[:kps1.size]

    def SearchByProjection(self, x3Dc, valid1, obs1, mp_desc1, octave1, angle1, kps2, desc2, u_right2, occupied2, cam, mbf,
                           bounds, scale_factors, th: float, direction: int = 0, orb_dist: int = TH_HIGH, depth_test: bool = True):
        """SearchByProjection(CurrentFrame, LastFrame, th, bMono), src/ORBmatcher.cc:1328-1470, with the boundary of
        include/orbx.h (camera-frame points in, assignment per current keypoint out).  Returns (nmatches, match21)."""
        x3Dc = np.ascontiguousarray(x3Dc, dtype=np.float32)
        valid1, obs1, occupied2 = (np.ascontiguousarray(a, dtype=np.uint8) for a in (valid1, obs1, occupied2))
        mp_desc1, desc2 = (np.ascontiguousarray(a, dtype=np.uint8) for a in (mp_desc1, desc2))
        octave1 = np.ascontiguousarray(octave1, dtype=np.int32)
        angle1 = np.ascontiguousarray(angle1, dtype=np.float32)
        kps2 = np.ascontiguousarray(kps2, dtype=KP_DTYPE)
        ur = None if u_right2 is None else np.ascontiguousarray(u_right2, dtype=np.float32)
        cam = np.asarray(cam, dtype=np.float32)
        b = np.asarray(bounds, dtype=np.float32)
        sf = np.ascontiguousarray(scale_factors, dtype=np.float32)
        n1, n2 = valid1.size, kps2.size
        m21 = np.empty(max(n2, 1), np.int32)
        n = C.c_int(0)
        _check(lib().orbx_search_by_projection_ex(n1, _p(x3Dc), _p(valid1), _p(obs1), _p(mp_desc1), _p(octave1), _p(angle1), _p(kps2),
                                                  n2, _p(desc2), _p(ur) if ur is not None else None, _p(occupied2), _p(cam),
                                                  C.c_float(mbf), _p(b), _p(sf), sf.size, C.c_float(th), int(direction),
                                                  int(self.mbCheckOrientation), int(orb_dist), 0 if depth_test else 1, _p(m21),
                                                  C.byref(n)))
        return n.value, m21[:n2]

    def SearchByProjectionMap(self, proj, view_cos, level1, valid1, obs1, mp_desc1, kps2, desc2, u_right2, occupied2, bounds,
                              scale_factors, th: float):
        """SearchByProjection(F, vpMapPoints, th), src/ORBmatcher.cc:45-129 (Tracking::SearchLocalPoints).
        Returns (nmatches, match21)."""
        proj = np.ascontiguousarray(proj, dtype=np.float32)
        view_cos = np.ascontiguousarray(view_cos, dtype=np.float32)
        level1 = np.ascontiguousarray(level1, dtype=np.int32)
        valid1, obs1, occupied2 = (np.ascontiguousarray(a, dtype=np.uint8) for a in (valid1, obs1, occupied2))
        mp_desc1, desc2 = (np.ascontiguousarray(a, dtype=np.uint8) for a in (mp_desc1, desc2))
        kps2 = np.ascontiguousarray(kps2, dtype=KP_DTYPE)
        ur = None if u_right2 is None else np.ascontiguousarray(u_right2, dtype=np.float32)
        b = np.asarray(bounds, dtype=np.float32)
        sf = np.ascontiguousarray(scale_factors, dtype=np.float32)
        n1, n2 = valid1.size, kps2.size
        m21 = np.empty(max(n2, 1), np.int32)
        n = C.c_int(0)
        _check(lib().orbx_search_by_projection_map(n1, _p(proj), _p(view_cos), _p(level1), _p(valid1), _p(obs1), _p(mp_desc1),
                                                   _p(kps2), n2, _p(desc2), _p(ur) if ur is not None else None, _p(occupied2),
                                                   _p(b), _p(sf), sf.size, C.c_float(th), C.c_float(self.mfNNratio), _p(m21),
                                                   C.byref(n)))
        return n.value, m21[:n2]

    def SearchByProjectionFrame(self, x3Dc, valid1, obs1, mp_desc1, octave1, angle1, frame2: "DeviceFrame", occupied2, cam, mbf, scale_factors,
                                th: float, direction: int = 0, orb_dist: int = TH_HIGH, depth_test: bool = True):
        """SearchByProjection against a frame that already lives on the device (orbx_search_by_projection_frame)."""
        x3Dc = np.ascontiguousarray(x3Dc, dtype=np.float32)
        valid1, obs1, occupied2 = (np.ascontiguousarray(a, dtype=np.uint8) for a in (valid1, obs1, occupied2))
        mp_desc1 = np.ascontiguousarray(mp_desc1, dtype=np.uint8)
        octave1 = np.ascontiguousarray(octave1, dtype=np.int32)
        angle1 = np.ascontiguousarray(angle1, dtype=np.float32)
        cam = np.asarray(cam, dtype=np.float32)
        sf = np.ascontiguousarray(scale_factors, dtype=np.float32)
        n2 = frame2.n
        m21 = np.empty(max(n2, 1), np.int32)
        n = C.c_int(0)
        _check(lib().orbx_search_by_projection_frame(valid1.size, _p(x3Dc), _p(valid1), _p(obs1), _p(mp_desc1), _p(octave1), _p(angle1),
                                                     frame2._h, _p(occupied2), _p(cam), C.c_float(mbf), _p(sf), sf.size, C.c_float(th),
                                                     int(direction), int(self.mbCheckOrientation), int(orb_dist), 0 if depth_test else 1,
                                                     _p(m21), C.byref(n)))
        return n.value, m21[:n2]

    def SearchByProjectionMapFrame(self, proj, view_cos, level1, valid1, obs1, mp_desc1, frame2: "DeviceFrame", occupied2, scale_factors, th: float):
        """SearchByProjection(F, vpMapPoints, th) against a device-resident frame (orbx_search_by_projection_map_frame)."""
        proj = np.ascontiguousarray(proj, dtype=np.float32)
        view_cos = np.ascontiguousarray(view_cos, dtype=np.float32)
        level1 = np.ascontiguousarray(level1, dtype=np.int32)
        valid1, obs1, occupied2 = (np.ascontiguousarray(a, dtype=np.uint8) for a in (valid1, obs1, occupied2))
        mp_desc1 = np.ascontiguousarray(mp_desc1, dtype=np.uint8)
        sf = np.ascontiguousarray(scale_factors, dtype=np.float32)
        n2 = frame2.n
        m21 = np.empty(max(n2, 1), np.int32)
        n = C.c_int(0)
        _check(lib().orbx_search_by_projection_map_frame(valid1.size, _p(proj), _p(view_cos), _p(level1), _p(valid1), _p(obs1), _p(mp_desc1),
                                                         frame2._h, _p(occupied2), _p(sf), sf.size, C.c_float(th), C.c_float(self.mfNNratio),
                                                         _p(m21), C.byref(n)))
        return n.value, m21[:n2]

    def SearchByProjectionSim3(self, uvr, level1, valid1, mp_desc1, kps2, desc2, occupied2, bounds, th_dist: int = TH_LOW):
        """SearchByProjection(pKF, Scw, vpPoints, vpMatched, th), src/ORBmatcher.cc:290-403, from the radius search on.
        Returns (nmatches, match21)."""
        uvr = np.ascontiguousarray(uvr, dtype=np.float32)
        level1 = np.ascontiguousarray(level1, dtype=np.int32)
        valid1, occupied2 = (np.ascontiguousarray(a, dtype=np.uint8) for a in (valid1, occupied2))
        mp_desc1, desc2 = (np.ascontiguousarray(a, dtype=np.uint8) for a in (mp_desc1, desc2))
        kps2 = np.ascontiguousarray(kps2, dtype=KP_DTYPE)
        b = np.asarray(bounds, dtype=np.float32)
        n1, n2 = valid1.size, kps2.size
        m21 = np.empty(max(n2, 1), np.int32)
        n = C.c_int(0)
        _check(lib().orbx_search_by_projection_sim3(n1, _p(uvr), _p(level1), _p(valid1), _p(mp_desc1), _p(kps2), n2, _p(desc2),
                                                    _p(occupied2), _p(b), int(th_dist), _p(m21), C.byref(n)))
        return n.value, m21[:n2]

    def SearchByBoW(self, start1, idx1, start2, idx2, valid1, desc1, angle1, valid2, desc2, angle2, th_dist: int = TH_LOW):
        """SearchByBoW(pKF, F, ...) (src/ORBmatcher.cc:159-288; th_dist = TH_LOW, valid2 all ones) and
        SearchByBoW(pKF1, pKF2, ...) (:522-655; th_dist = TH_LOW - 1) over the common vocabulary nodes start*/idx*.
        Returns (nmatches, match12, match21)."""
        start1, start2 = (np.ascontiguousarray(a, dtype=np.int32) for a in (start1, start2))
        idx1, idx2 = (np.ascontiguousarray(a, dtype=np.uint32) for a in (idx1, idx2))
        valid1, valid2 = (np.ascontiguousarray(a, dtype=np.uint8) for a in (valid1, valid2))
        desc1, desc2 = (np.ascontiguousarray(a, dtype=np.uint8) for a in (desc1, desc2))
        angle1, angle2 = (np.ascontiguousarray(a, dtype=np.float32) for a in (angle1, angle2))
        n_nodes, n1, n2 = max(start1.size - 1, 0), valid1.size, valid2.size
        m12, m21 = np.empty(max(n1, 1), np.int32), np.empty(max(n2, 1), np.int32)
        n = C.c_int(0)
        _check(lib().orbx_search_by_bow(n_nodes, _p(start1), _p(idx1), _p(start2), _p(idx2), n1, _p(valid1), _p(desc1), _p(angle1),
                                        n2, _p(valid2), _p(desc2), _p(angle2), int(th_dist), C.c_float(self.mfNNratio),
                                        int(self.mbCheckOrientation), _p(m12), _p(m21), C.byref(n)))
        return n.value, m12[:n1], m21[:n2]

    def FuseMatch(self, uvr, ur1, level1, valid1, mp_desc1, kps2, desc2, u_right2, inv_level_sigma2, bounds):
        """The candidate loop of ORBmatcher::Fuse (src/ORBmatcher.cc:886-945 with the chi-square gate, ur1 given; :1049-1075 without,
        ur1 = None).  Returns (best_idx, best_dist) per map point."""
        uvr = np.ascontiguousarray(uvr, dtype=np.float32)
        ur1 = None if ur1 is None else np.ascontiguousarray(ur1, dtype=np.float32)
        level1 = np.ascontiguousarray(level1, dtype=np.int32)
        valid1, mp_desc1, desc2 = (np.ascontiguousarray(a, dtype=np.uint8) for a in (valid1, mp_desc1, desc2))
        kps2 = np.ascontiguousarray(kps2, dtype=KP_DTYPE)
        u_right2 = None if u_right2 is None else np.ascontiguousarray(u_right2, dtype=np.float32)
        sg = np.ascontiguousarray(inv_level_sigma2, dtype=np.float32)
        b = np.asarray(bounds, dtype=np.float32)
        n1 = valid1.size
        idx, dist = np.empty(max(n1, 1), np.int32), np.empty(max(n1, 1), np.int32)
        _check(lib().orbx_fuse_match(n1, _p(uvr), _p(ur1), _p(level1), _p(valid1), _p(mp_desc1), _p(kps2), kps2.size, _p(desc2),
                                     _p(u_right2), _p(sg), sg.size, _p(b), _p(idx), _p(dist)))
        return idx[:n1], dist[:n1]

    def SearchForTriangulation(self, start1, idx1, start2, idx2, kps1, desc1, valid1, stereo1, kps2, desc2, valid2, stereo2, F12, ex, ey,
                               scale_factors2, level_sigma2_2):
        """SearchForTriangulation(pKF1, pKF2, F12, vMatchedPairs, bOnlyStereo), src/ORBmatcher.cc:657-823.
        Returns (nmatches, match12); vMatchedPairs = [(i, match12[i]) for i where match12[i] >= 0]."""
        start1, start2 = (np.ascontiguousarray(a, dtype=np.int32) for a in (start1, start2))
        idx1, idx2 = (np.ascontiguousarray(a, dtype=np.uint32) for a in (idx1, idx2))
        kps1, kps2 = (np.ascontiguousarray(a, dtype=KP_DTYPE) for a in (kps1, kps2))
        desc1, desc2, valid1, valid2, stereo1, stereo2 = (np.ascontiguousarray(a, dtype=np.uint8)
                                                          for a in (desc1, desc2, valid1, valid2, stereo1, stereo2))
        F = np.ascontiguousarray(F12, dtype=np.float32).reshape(9)
        sf, sg = (np.ascontiguousarray(a, dtype=np.float32) for a in (scale_factors2, level_sigma2_2))
        n1, n2 = kps1.size, kps2.size
        m12 = np.empty(max(n1, 1), np.int32)
        n = C.c_int(0)
        _check(lib().orbx_search_for_triangulation(max(start1.size - 1, 0), _p(start1), _p(idx1), _p(start2), _p(idx2), n1, _p(kps1),
                                                   _p(desc1), _p(valid1), _p(stereo1), n2, _p(kps2), _p(desc2), _p(valid2), _p(stereo2),
                                                   _p(F), C.c_float(ex), C.c_float(ey), _p(sf), _p(sg), sf.size,
                                                   int(self.mbCheckOrientation), _p(m12), C.byref(n)))
        return n.value, m12[:n1]

    def match_bruteforce(self, q, d, th: int = TH_LOW):
        """best / second-best over the whole database (inner loop of SearchByBoW, src/ORBmatcher.cc:187-232):
        returns (idx, d1, d2, match) int32 [nq]."""
        q = np.ascontiguousarray(q, dtype=np.uint8)
        d = np.ascontiguousarray(d, dtype=np.uint8)
        nq, nd = q.shape[0], d.shape[0]
        idx, d1, d2, m = (np.empty(max(nq, 1), np.int32) for _ in range(4))
        _check(lib().orbx_match_bruteforce(_p(q), nq, _p(d), nd, int(th), C.c_float(self.mfNNratio), _p(idx), _p(d1), _p(d2), _p(m)))
        return idx[:nq], d1[:nq], d2[:nq], m[:nq]

    def match_bruteforce_device(self, d_q: int, nq: int, d_d: int, nd: int, d_idx: int, d_d1: int, d_d2: int, d_match: int,
                                th: int = TH_LOW, stream: int = 0):
        _check(lib().orbx_match_bruteforce_device(_p(d_q), nq, _p(d_d), nd, int(th), C.c_float(self.mfNNratio), _p(d_idx),
                                                  _p(d_d1), _p(d_d2), _p(d_match), _p(stream) if stream else None))

    def match_bruteforce_noalloc(self, d_q: int, nq: int, d_d: int, nd: int, d_idx: int, d_d1: int, d_d2: int, d_match: int,
                                 th: int = TH_LOW, stream: int = 0):
        """orbx_match_bruteforce_noalloc: match_bruteforce_device with its scratch in a static device array of liborbx.so —
        allocates nothing; nq <= NOALLOC_MAX_QUERIES; calls that overlap on the GPU race on the shared scratch."""
        _check(lib().orbx_match_bruteforce_noalloc(_p(d_q), nq, _p(d_d), nd, int(th), C.c_float(self.mfNNratio), _p(d_idx),
                                                   _p(d_d1), _p(d_d2), _p(d_match), _p(stream) if stream else None))

    def match_allpairs_device(self, d_q: int, kq: int, d_db: int, kd: int, n: int, d_score: int, th: int = TH_LOW, stream: int = 0):
        """score[i, j] = accepted matches of query keyframe i in database keyframe j (device buffers, asynchronous)."""
        _check(lib().orbx_match_allpairs_device(_p(d_q), kq, _p(d_db), kd, n, int(th), C.c_float(self.mfNNratio), _p(d_score),
                                                _p(stream) if stream else None))


NOALLOC_MAX_QUERIES = 8192  # ORBX_NOALLOC_MAX_QUERIES of include/orbx.h
GRAY_BGR, GRAY_CV2_COEFFS = 1, 2


def cvtColorGray(img, bgr: bool = False, cv2_coeffs: bool = False):
    """cvtColor(img, CV_RGB2GRAY / BGR2GRAY / RGBA2GRAY / BGRA2GRAY) as Tracking::GrabImage* does (src/Tracking.cc:221-246):
    img uint8 [h, w, 3 or 4] -> uint8 [h, w]."""
    img = np.ascontiguousarray(img, dtype=np.uint8)
    h, w, ch = img.shape
    out = np.empty((h, w), np.uint8)
    _check(lib().orbx_cvt_color_gray(_p(img), C.c_longlong(w * ch), w, h, ch, (GRAY_BGR if bgr else 0) | (GRAY_CV2_COEFFS if cv2_coeffs else 0),
                                     _p(out), C.c_longlong(w)))
    return out


def ComputeDistinctiveDescriptors(start, desc):
    """MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:246-312) for a batch of map points: rows start[p]..start[p+1] of
    desc are point p's observed descriptors.  Returns (best_idx, best_median) int32 [n_points]."""
    start = np.ascontiguousarray(start, dtype=np.int32)
    desc = np.ascontiguousarray(desc, dtype=np.uint8)
    n = max(start.size - 1, 0)
    idx, med = np.empty(max(n, 1), np.int32), np.empty(max(n, 1), np.int32)
    _check(lib().orbx_distinctive_descriptors(n, _p(start), _p(desc), _p(idx), _p(med)))
    return idx[:n], med[:n]


class ORBVocabulary:
    """ORB_SLAM2::ORBVocabulary = DBoW2::TemplatedVocabulary<FORB::TDescriptor, FORB> (reference include/ORBVocabulary.h:28-31):
    the tree lives in device memory, transform() is Frame::ComputeBoW's call (src/Frame.cc:394-401)."""

    def __init__(self, k, L, scoring, weighting, parent, is_leaf, desc, weight, _handle=None):
        self._h = C.c_void_p()
        if _handle is not None:
            self._h = _handle
            return
        parent = np.ascontiguousarray(parent, dtype=np.int32)
        is_leaf = np.ascontiguousarray(is_leaf, dtype=np.uint8)
        desc = np.ascontiguousarray(desc, dtype=np.uint8)
        weight = np.ascontiguousarray(weight, dtype=np.float64)
        _check(lib().orbx_vocabulary_create(int(k), int(L), int(scoring), int(weighting), parent.size, _p(parent), _p(is_leaf), _p(desc),
                                            _p(weight), C.byref(self._h)))

    @classmethod
    def loadFromTextFile(cls, path: str):
        h = C.c_void_p()
        _check(lib().orbx_vocabulary_load_text(str(path).encode(), C.byref(h)))
        return cls(0, 0, 0, 0, None, None, None, None, _handle=h)

    def info(self):
        v = [C.c_int(0) for _ in range(6)]
        _check(lib().orbx_vocabulary_info(self._h, *[C.byref(x) for x in v]))
        return dict(zip(("k", "L", "scoring", "weighting", "n_nodes", "n_words"), (x.value for x in v)))

    def transform_features(self, desc, levelsup: int = 4):
        """(word_id, weight, node_id) per descriptor: transform(feature, id, w, &nid, levelsup), TemplatedVocabulary.h:1218-1261."""
        desc = np.ascontiguousarray(desc, dtype=np.uint8)
        n = desc.shape[0]
        word, node, w = np.empty(max(n, 1), np.uint32), np.empty(max(n, 1), np.uint32), np.empty(max(n, 1), np.float64)
        _check(lib().orbx_vocabulary_transform(self._h, _p(desc), n, int(levelsup), _p(word), _p(w), _p(node)))
        return word[:n], w[:n], node[:n]

    def transform(self, desc, levelsup: int = 4):
        """transform(features, BowVector&, FeatureVector&, levelsup), :1127-1194: ({word: value}, {node: [feature indices]}).
        The descent runs on the device; the two std::map fills and the normalisation are sequential double arithmetic in feature
        / word order and stay on the host (same code as include/ORBVocabulary.h)."""
        info = self.info()
        word, w, node = self.transform_features(desc, levelsup)
        bow, fv, nid = {}, {}, 0
        tf = info["weighting"] in (0, 1)
        for i in range(word.size):
            if node[i] != 0xFFFFFFFF:
                nid = int(node[i])
            if w[i] > 0:
                if tf:
                    bow[int(word[i])] = bow.get(int(word[i]), 0.0) + float(w[i])
                else:
                    bow.setdefault(int(word[i]), float(w[i]))
                fv.setdefault(nid, []).append(i)
        must, l1 = info["scoring"] != 5, info["scoring"] != 1
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

    def __del__(self):
        try:
            if self._h:
                lib().orbx_vocabulary_destroy(self._h)
                self._h = C.c_void_p()
        except Exception:
            pass


def ComputeStereoMatches(exL: ORBextractor, exR: ORBextractor, kpsL, descL, kpsR, descR, bf: float, fx: float):
    """Frame::ComputeStereoMatches (src/Frame.cc:465-639) on the pyramids exL / exR hold from their last call.
    Returns (mvuRight, mvDepth) float32 [nL], -1 where there is no match."""
    kpsL = np.ascontiguousarray(kpsL, dtype=KP_DTYPE)
    kpsR = np.ascontiguousarray(kpsR, dtype=KP_DTYPE)
    descL = np.ascontiguousarray(descL, dtype=np.uint8)
    descR = np.ascontiguousarray(descR, dtype=np.uint8)
    u = np.empty(max(kpsL.size, 1), np.float32)
    z = np.empty(max(kpsL.size, 1), np.float32)
    _check(lib().orbx_stereo_match(exL._h, exR._h, _p(kpsL), kpsL.size, _p(descL), _p(kpsR), kpsR.size, _p(descR),
                                   C.c_float(bf), C.c_float(np.float32(bf) / np.float32(fx)), _p(u), _p(z)))
    return u[:kpsL.size], z[:kpsL.size]


def ComputeStereoMatchesLast(exL: ORBextractor, exR: ORBextractor, n_left: int, bf: float, fx: float):
    """orbx_stereo_match_last: Frame::ComputeStereoMatches on the keypoints / descriptors the two extractors produced in their
    last single-frame call, still on the device (nothing is uploaded).  n_left = number of left keypoints of that call."""
    u = np.empty(max(n_left, 1), np.float32)
    z = np.empty(max(n_left, 1), np.float32)
    _check(lib().orbx_stereo_match_last(exL._h, exR._h, C.c_float(bf), C.c_float(np.float32(bf) / np.float32(fx)), _p(u), _p(z), int(n_left)))
    return u[:n_left], z[:n_left]


def UndistortKeyPoints(kps: np.ndarray, K, dist) -> np.ndarray:
    """Frame::UndistortKeyPoints (reference src/Frame.cc:403-433): K = (fx, fy, cx, cy), dist = k1 k2 p1 p2 [k3]."""
    kps = np.ascontiguousarray(kps, dtype=KP_DTYPE)
    Kf = np.asarray(K, dtype=np.float32)
    d = np.asarray(dist, dtype=np.float32).reshape(-1)
    out = np.empty_like(kps)
    _check(lib().orbx_undistort_keypoints(_p(kps), kps.size, _p(Kf), _p(d) if d.size else None, int(d.size), _p(out)))
    return out


class DeviceFrame:
    """orbx_frame: the undistorted keypoints, descriptors, right coordinates and the 64 x 48 grid (Frame::AssignFeaturesToGrid,
    reference src/Frame.cc:229-244) of one frame, resident on the device between matcher calls."""

    def __init__(self, kps_un, desc, bounds, u_right=None):
        kps_un = np.ascontiguousarray(kps_un, dtype=KP_DTYPE)
        desc = np.ascontiguousarray(desc, dtype=np.uint8)
        ur = None if u_right is None else np.ascontiguousarray(u_right, dtype=np.float32)
        b = np.asarray(bounds, dtype=np.float32)
        self._h = C.c_void_p()
        self.n = int(kps_un.size)
        _check(lib().orbx_frame_create(_p(kps_un), self.n, _p(desc), _p(ur) if ur is not None else None, _p(b), C.byref(self._h)))

    def grid(self):
        """(cell_start [64*48 + 1], sorted [n]): cell (ix, iy) holds sorted[cell_start[ix*48 + iy] : cell_start[ix*48 + iy + 1]]."""
        cs, so = np.empty(64 * 48 + 1, np.int32), np.empty(max(self.n, 1), np.int32)
        lib().orbx_frame_grid_host.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        _check(lib().orbx_frame_grid_host(self._h, _p(cs), _p(so)))
        return cs, so[:self.n]

    def __del__(self):
        h = getattr(self, "_h", None)
        if h and _lib is not None:
            _lib.orbx_frame_destroy.argtypes = [C.c_void_p]
            _lib.orbx_frame_destroy.restype = None
            _lib.orbx_frame_destroy(h)
            self._h = None
