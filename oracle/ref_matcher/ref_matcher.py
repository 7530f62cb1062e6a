"""TEST INFRASTRUCTURE: ctypes access to oracle/_ref/libref_matcher.so — the reference's own src/ORBmatcher.cc and src/Frame.cc
compiled from /root/reference by oracle/ref_matcher/Makefile (stand-ins for the rest of the SLAM system: ref_stubs.h, a
test-only OpenCV shim).  Only tests/ and tools/gen_matcher_golden.py use it; argument conventions follow oracle_lib's."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(_HERE, "..", "_ref", "libref_matcher.so")
REF = "/root/reference/src/ORBmatcher.cc"
KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"), ("octave", "<i4"),
                     ("class_id", "<i4")])


def build():
    """Compiles the reference sources where they lie; returns the .so path or None when /root/reference is absent (GPU box)."""
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
    return _lib


def available():
    return lib() is not None


def _p(a):
    return None if a is None else C.c_void_p(a.ctypes.data)


def _u8(a):
    return np.ascontiguousarray(a, dtype=np.uint8)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _kp(a):
    return np.ascontiguousarray(a, dtype=KP_DTYPE)


def _wh(bounds):
    assert bounds[0] == 0 and bounds[2] == 0, "the reference Frame of the harness has an undistorted camera: bounds (0, w, 0, h)"
    return np.array([int(bounds[1]), int(bounds[3])], np.int32)


def descriptor_distance(a, b):
    return int(lib().refm_descriptor_distance(_p(_u8(a)), _p(_u8(b))))


def features_in_area(keys, bounds, x, y, r, min_level=-1, max_level=-1):
    keys = _kp(keys)
    out = np.empty(max(keys.size, 1), np.int32)
    n = lib().refm_features_in_area(_p(keys), keys.size, _p(_wh(bounds)), C.c_float(x), C.c_float(y), C.c_float(r), int(min_level),
                                    int(max_level), _p(out))
    return out[:n].copy()


def search_for_initialization(k1, d1, k2, d2, bounds, prev_xy, window=100, nnratio=0.9, check_ori=True):
    k1, k2, d1, d2 = _kp(k1), _kp(k2), _u8(d1), _u8(d2)
    prev = _f32(prev_xy).copy()
    m12 = np.empty(max(k1.size, 1), np.int32)
    n = lib().refm_search_for_initialization(_p(k1), k1.size, _p(d1), _p(k2), k2.size, _p(d2), _p(_wh(bounds)), _p(prev), _p(m12),
                                             int(window), C.c_float(nnratio), int(check_ori))
    return n, m12[:k1.size], prev


def search_by_projection_frame(x3Dc, valid1, obs1, mp_desc1, k1, k2, d2, u_right2, occupied2, cam, mbf, bounds, scale_factors, th,
                               direction=0, check_ori=True):
    """direction 0: monocular (no forward / backward); 1: forward; 2: backward — set through LastFrame's pose against
    CurrentFrame.mb = mbf / fx (src/ORBmatcher.cc:1354-1357)."""
    x3, valid1, obs1, occupied2 = _f32(x3Dc), _u8(valid1), _u8(obs1), _u8(occupied2)
    mp, k1, k2, d2 = _u8(mp_desc1), _kp(k1), _kp(k2), _u8(d2)
    ur = None if u_right2 is None else _f32(u_right2)
    cam, sf = _f32(cam), _f32(scale_factors)
    mb = float(mbf) / float(cam[0])
    tlcz = {0: 0.0, 1: 2.0 * mb + 1.0, 2: -2.0 * mb - 1.0}[direction]
    m21 = np.empty(max(k2.size, 1), np.int32)
    n = lib().refm_search_by_projection_frame(valid1.size, _p(x3), _p(valid1), _p(obs1), _p(mp), _p(k1), _p(k2), k2.size, _p(d2), _p(ur),
                                              _p(occupied2), _p(_wh(bounds)), _p(cam), C.c_float(mbf), _p(sf), sf.size, C.c_float(th),
                                              C.c_float(tlcz), int(direction == 0), int(check_ori), _p(m21))
    return n, m21[:k2.size]


def search_by_projection_map(proj, view_cos, level1, valid1, obs1, mp_desc1, k2, d2, u_right2, occupied2, bounds, scale_factors, th,
                             nnratio=0.8):
    proj, view_cos, level1 = _f32(proj), _f32(view_cos), _i32(level1)
    valid1, obs1, occupied2, mp, k2, d2 = _u8(valid1), _u8(obs1), _u8(occupied2), _u8(mp_desc1), _kp(k2), _u8(d2)
    ur = None if u_right2 is None else _f32(u_right2)
    sf = _f32(scale_factors)
    m21 = np.empty(max(k2.size, 1), np.int32)
    n = lib().refm_search_by_projection_map(valid1.size, _p(proj), _p(view_cos), _p(level1), _p(valid1), _p(obs1), _p(mp), _p(k2), k2.size,
                                            _p(d2), _p(ur), _p(occupied2), _p(_wh(bounds)), _p(sf), sf.size, C.c_float(th),
                                            C.c_float(nnratio), _p(m21))
    return n, m21[:k2.size]


def search_by_bow_frame(start1, idx1, start2, idx2, valid1, k1, d1, k2, d2, bounds, nnratio=0.7, check_ori=True):
    s1, s2, i1, i2 = _i32(start1), _i32(start2), np.ascontiguousarray(idx1, np.uint32), np.ascontiguousarray(idx2, np.uint32)
    valid1, k1, d1, k2, d2 = _u8(valid1), _kp(k1), _u8(d1), _kp(k2), _u8(d2)
    m21 = np.empty(max(k2.size, 1), np.int32)
    n = lib().refm_search_by_bow_frame(s1.size - 1, _p(s1), _p(i1), _p(s2), _p(i2), k1.size, _p(valid1), _p(k1), _p(d1), k2.size, _p(k2),
                                       _p(d2), _p(_wh(bounds)), C.c_float(nnratio), int(check_ori), _p(m21))
    return n, m21[:k2.size]


def search_by_bow_keyframes(start1, idx1, start2, idx2, valid1, k1, d1, valid2, k2, d2, bounds, nnratio=0.75, check_ori=True):
    s1, s2, i1, i2 = _i32(start1), _i32(start2), np.ascontiguousarray(idx1, np.uint32), np.ascontiguousarray(idx2, np.uint32)
    valid1, k1, d1, valid2, k2, d2 = _u8(valid1), _kp(k1), _u8(d1), _u8(valid2), _kp(k2), _u8(d2)
    m12 = np.empty(max(k1.size, 1), np.int32)
    n = lib().refm_search_by_bow_keyframes(s1.size - 1, _p(s1), _p(i1), _p(s2), _p(i2), k1.size, _p(valid1), _p(k1), _p(d1), k2.size,
                                           _p(valid2), _p(k2), _p(d2), _p(_wh(bounds)), C.c_float(nnratio), int(check_ori), _p(m12))
    return n, m12[:k1.size]


def search_for_triangulation(start1, idx1, start2, idx2, k1, d1, valid1, stereo1, k2, d2, valid2, stereo2, F12, Cw, cam, bounds,
                             scale_factors, only_stereo=False, check_ori=True):
    """Cw: camera centre of keyframe 1 in keyframe 2's frame; the reference derives the epipole from it (src/ORBmatcher.cc:664-670)."""
    s1, s2, i1, i2 = _i32(start1), _i32(start2), np.ascontiguousarray(idx1, np.uint32), np.ascontiguousarray(idx2, np.uint32)
    k1, d1, valid1, k2, d2, valid2 = _kp(k1), _u8(d1), _u8(valid1), _kp(k2), _u8(d2), _u8(valid2)
    ur1 = np.where(_u8(stereo1) != 0, 10.0, -1.0).astype(np.float32)
    ur2 = np.where(_u8(stereo2) != 0, 10.0, -1.0).astype(np.float32)
    F, Cw, cam, sf = _f32(F12).reshape(9), _f32(Cw), _f32(cam), _f32(scale_factors)
    m12 = np.empty(max(k1.size, 1), np.int32)
    n = lib().refm_search_for_triangulation(max(s1.size - 1, 0), _p(s1), _p(i1), _p(s2), _p(i2), k1.size, _p(k1), _p(d1), _p(valid1),
                                            _p(ur1), k2.size, _p(k2), _p(d2), _p(valid2), _p(ur2), _p(_wh(bounds)), _p(cam), _p(sf),
                                            sf.size, _p(F), _p(Cw), int(only_stereo), int(check_ori), _p(m12))
    return n, m12[:k1.size]


def search_by_projection_sim3(x3D, normal, minmax, level1, bad1, found1, mp_desc1, k2, d2, occupied2, cam, bounds, scale_factors, th):
    x3, nr, mm, level1 = _f32(x3D), _f32(normal), _f32(minmax), _i32(level1)
    bad1, found1, mp, k2, d2, occupied2 = _u8(bad1), _u8(found1), _u8(mp_desc1), _kp(k2), _u8(d2), _u8(occupied2)
    cam, sf = _f32(cam), _f32(scale_factors)
    m21 = np.empty(max(k2.size, 1), np.int32)
    n = lib().refm_search_by_projection_sim3(level1.size, _p(x3), _p(nr), _p(mm), _p(level1), _p(bad1), _p(found1), _p(mp), _p(k2), k2.size,
                                             _p(d2), _p(occupied2), _p(_wh(bounds)), _p(cam), _p(sf), sf.size, int(th), _p(m21))
    return n, m21[:k2.size]


def fuse(x3D, normal, minmax, level1, bad1, obs1, mp_desc1, k2, d2, u_right2, kf_point_obs, cam, mbf, bounds, scale_factors, th):
    """ORBmatcher::Fuse(pKF, vpMapPoints, th).  Returns (nFused, ops) with ops an int64 [n, 3] log of the map surgery:
    (0, point, idx) AddObservation, (1, point, idx) AddMapPoint, (2, a, b) a->Replace(b); the keyframe's own points are
    numbered len(level1) + feature index."""
    x3, nr, mm, level1 = _f32(x3D), _f32(normal), _f32(minmax), _i32(level1)
    bad1, obs1, mp, k2, d2 = _u8(bad1), _i32(obs1), _u8(mp_desc1), _kp(k2), _u8(d2)
    ur = np.full(k2.size, -1.0, np.float32) if u_right2 is None else _f32(u_right2)
    kpo, cam, sf = _i32(kf_point_obs), _f32(cam), _f32(scale_factors)
    cap = 3 * (level1.size + 8)
    ops = np.zeros((cap, 3), np.int64)
    nops = C.c_int(0)
    n = lib().refm_fuse(level1.size, _p(x3), _p(nr), _p(mm), _p(level1), _p(bad1), _p(obs1), _p(mp), _p(k2), k2.size, _p(d2), _p(ur), _p(kpo),
                        _p(_wh(bounds)), _p(cam), C.c_float(mbf), _p(sf), sf.size, C.c_float(th), _p(ops), cap, C.byref(nops))
    return n, ops[:nops.value].copy()


def stereo_matches(kL, dL, kR, dR, levelsL, levelsR, wh, cam, bf, scale_factors):
    """Frame::ComputeStereoMatches through the reference's stereo constructor; levels*: lists of 2-D uint8 arrays (interior)."""
    kL, dL, kR, dR = _kp(kL), _u8(dL), _kp(kR), _u8(dR)
    lv_l = [np.ascontiguousarray(a, np.uint8) for a in levelsL]
    lv_r = [np.ascontiguousarray(a, np.uint8) for a in levelsR]
    nl = len(lv_l)
    arr_l = (C.c_void_p * nl)(*[a.ctypes.data for a in lv_l])
    arr_r = (C.c_void_p * nl)(*[a.ctypes.data for a in lv_r])
    lwh = np.array([[a.shape[1], a.shape[0]] for a in lv_l], np.int32)
    cam, sf = _f32(cam), _f32(scale_factors)
    ur, depth = np.empty(max(kL.size, 1), np.float32), np.empty(max(kL.size, 1), np.float32)
    lib().refm_stereo_matches(_p(kL), kL.size, _p(dL), _p(kR), kR.size, _p(dR), arr_l, arr_r, _p(lwh), _p(np.array(wh, np.int32)), _p(cam),
                              C.c_float(bf), _p(sf), sf.size, _p(ur), _p(depth))
    return ur[:kL.size], depth[:kL.size]


def search_by_projection_reloc(x3Dc, has1, bad1, found1, minmax, level1, mp_desc1, k1, k2, d2, occupied2, cam, bounds, scale_factors, th,
                               orb_dist=100, check_ori=True):
    """ORBmatcher::SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist) with CurrentFrame at the identity pose."""
    x3, has1, bad1, found1, occupied2 = _f32(x3Dc), _u8(has1), _u8(bad1), _u8(found1), _u8(occupied2)
    mm, level1, mp, k1, k2, d2 = _f32(minmax), _i32(level1), _u8(mp_desc1), _kp(k1), _kp(k2), _u8(d2)
    cam, sf = _f32(cam), _f32(scale_factors)
    m21 = np.empty(max(k2.size, 1), np.int32)
    n = lib().refm_search_by_projection_reloc(has1.size, _p(x3), _p(has1), _p(bad1), _p(found1), _p(mm), _p(level1), _p(mp), _p(k1), _p(k2),
                                              k2.size, _p(d2), _p(occupied2), _p(_wh(bounds)), _p(cam), _p(sf), sf.size, C.c_float(th),
                                              int(orb_dist), int(check_ori), _p(m21))
    return n, m21[:k2.size]


def fuse_sim3(x3D, normal, minmax, level1, bad1, mp_desc1, entry_point, k2, d2, kf_point_obs, kf_bad, cam, bounds, scale_factors, scale, th):
    """ORBmatcher::Fuse(pKF, Scw, vpPoints, th, vpReplacePoint) with Scw = scale * [I | 0].  Returns (nFused, ops, replaced)."""
    x3, nr, mm, level1 = _f32(x3D), _f32(normal), _f32(minmax), _i32(level1)
    bad1, mp, ent, k2, d2 = _u8(bad1), _u8(mp_desc1), _i32(entry_point), _kp(k2), _u8(d2)
    kpo, kfb, cam, sf = _i32(kf_point_obs), _u8(kf_bad), _f32(cam), _f32(scale_factors)
    cap = 3 * (ent.size + 8)
    ops = np.zeros((cap, 3), np.int64)
    nops = C.c_int(0)
    rep = np.empty(max(ent.size, 1), np.int32)
    n = lib().refm_fuse_sim3(level1.size, _p(x3), _p(nr), _p(mm), _p(level1), _p(bad1), _p(mp), ent.size, _p(ent), _p(k2), k2.size, _p(d2),
                             _p(kpo), _p(kfb), _p(_wh(bounds)), _p(cam), _p(sf), sf.size, C.c_float(scale), C.c_float(th), _p(ops), cap,
                             C.byref(nops), _p(rep))
    return n, ops[:nops.value].copy(), rep[:ent.size]


def search_by_sim3(k1, d1, has1, bad1, x3D1, level1, k2, d2, has2, bad2, x3D2, level2, minmax, matched12, cam, bounds, scale_factors, s12, t12, th):
    """ORBmatcher::SearchBySim3 with both keyframes at the identity pose and R12 = I.  Returns (nFound, vpMatches12 as keyframe-2 indices)."""
    k1, d1, has1, bad1, x1, l1 = _kp(k1), _u8(d1), _u8(has1), _u8(bad1), _f32(x3D1), _i32(level1)
    k2, d2, has2, bad2, x2, l2 = _kp(k2), _u8(d2), _u8(has2), _u8(bad2), _f32(x3D2), _i32(level2)
    mm, m12, cam, sf, t = _f32(minmax), _i32(matched12), _f32(cam), _f32(scale_factors), _f32(t12)
    out = np.empty(max(k1.size, 1), np.int32)
    n = lib().refm_search_by_sim3(k1.size, _p(k1), _p(d1), _p(has1), _p(bad1), _p(x1), _p(l1), k2.size, _p(k2), _p(d2), _p(has2), _p(bad2),
                                  _p(x2), _p(l2), _p(mm), _p(m12), _p(_wh(bounds)), _p(cam), _p(sf), sf.size, C.c_float(s12), _p(t),
                                  C.c_float(th), _p(out))
    return n, out[:k1.size]
