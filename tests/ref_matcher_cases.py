"""Cases that pin oracle/orb_oracle.cpp to the REFERENCE's own object code (oracle/_ref/libref_matcher.so = /root/reference's
src/ORBmatcher.cc + src/Frame.cc compiled by oracle/ref_matcher/Makefile).  Every case is a pair of callables returning tuples of
arrays that must be equal: `oracle` runs the restatement, `reference` runs the compiled reference through the harness.
tests/test_oracle_ref_matcher.py compares them live (where the reference library exists) and against tests/golden/ref_matcher.npz
(made by tools/gen_matcher_golden.py) everywhere else."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from oracle import oracle_lib as O  # noqa: E402
from oracle.ref_matcher import ref_matcher as R  # noqa: E402
from orb_slam_cuda_b200 import synth  # noqa: E402

import test_bow as TB  # noqa: E402
import test_projection as TP  # noqa: E402
import test_triangulation as TT  # noqa: E402

F32 = np.float32


def _null(m):  # the oracle marks matches removed by the rotation check with -2, the reference sets them to NULL
    return np.where(m == -2, -1, m).astype(np.int32)


def _extract(kind, k, nfeat, shifted=False):
    img = synth.shifted_frame(kind, k) if shifted else synth.frame(kind, k)
    ox = O.Extractor(nfeat, 1.2, 8, 20, 7)
    kps, desc = ox.extract(img)
    return kps.copy(), desc.copy(), img.shape


def case_descriptor_distance():
    rng = np.random.default_rng(5)
    a, b = rng.integers(0, 256, (200, 32), dtype=np.uint8), rng.integers(0, 256, (200, 32), dtype=np.uint8)
    b[:50] = a[:50] ^ (1 << rng.integers(0, 8, (50, 32))).astype(np.uint8) * (rng.random((50, 32)) < 0.2)
    return (lambda: (np.array([O.descriptor_distance(x, y) for x, y in zip(a, b)], np.int32),),
            lambda: (np.array([R.descriptor_distance(x, y) for x, y in zip(a, b)], np.int32),))


def case_features_in_area():
    k, _, (h, w) = _extract("tum", 3, 1000)
    rng = np.random.default_rng(9)
    qs = [(float(rng.uniform(-20, w + 20)), float(rng.uniform(-20, h + 20)), float(rng.uniform(1, 120)), int(a), int(b))
          for a, b in [(-1, -1), (0, 2), (3, -1), (1, 1), (0, 7), (-1, 3), (2, 5), (-1, -1)] for _ in range(6)]
    bounds = (0.0, float(w), 0.0, float(h))

    def oracle():
        out = []
        for x, y, r, a, b in qs:
            out.append(np.array(O.features_in_area(k, bounds, x, y, r, a, b), np.int32))
        return (np.concatenate(out), np.array([o.size for o in out], np.int32))

    def reference():
        out = [R.features_in_area(k, bounds, x, y, r, a, b) for x, y, r, a, b in qs]
        return (np.concatenate(out), np.array([o.size for o in out], np.int32))
    return oracle, reference


def case_sfi(kind, k, nfeat, window, ratio, ori, random_desc=False):
    k1, d1, (h, w) = _extract(kind, k, nfeat)
    k2, d2, _ = _extract(kind, k, nfeat, shifted=True)
    if random_desc:  # few distinct descriptors: many equal distances, heavy competition for the same keypoints
        rng = np.random.default_rng(k)
        pool = rng.integers(0, 256, (12, 32), dtype=np.uint8)
        d1, d2 = pool[rng.integers(0, 12, k1.size)], pool[rng.integers(0, 12, k2.size)]
        d1, d2 = d1 ^ (rng.random(d1.shape) < 0.02).astype(np.uint8), d2 ^ (rng.random(d2.shape) < 0.02).astype(np.uint8)
    bounds = (0.0, float(w), 0.0, float(h))
    prev = np.stack([k1["x"], k1["y"]], 1).astype(F32)

    def two_rounds(fn):
        n1, m1, p1 = fn(k1, d1, k2, d2, bounds, prev, window, ratio, ori)
        n2, m2, p2 = fn(k1, d1, k2, d2, bounds, p1, max(window // 2, 5), ratio, ori)  # second call with the updated vbPrevMatched
        return (np.array([n1, n2], np.int32), m1.copy(), p1.view(np.uint32).copy(), m2.copy(), p2.view(np.uint32).copy())
    return (lambda: two_rounds(O.search_for_initialization)), (lambda: two_rounds(R.search_for_initialization))


def case_sbp_frame(seed, stereo, th, direction, ori):
    s = TP.scenario(seed, stereo)
    k1 = np.zeros(s["valid1"].size, O.KP_DTYPE)
    k1["octave"], k1["angle"], k1["x"], k1["y"] = s["octave1"], s["angle1"], 50.0, 50.0

    def oracle():
        n, m = O.search_by_projection_frame(s["x3Dc"], s["valid1"], s["obs1"], s["mp_desc1"], s["octave1"], s["angle1"], s["k2"], s["d2"],
                                            s["u_right2"], s["occupied2"], s["cam"], s["mbf"], s["bounds"], s["scale_factors"], th, direction, ori)
        return (np.array([n], np.int32), _null(m))

    def reference():
        n, m = R.search_by_projection_frame(s["x3Dc"], s["valid1"], s["obs1"], s["mp_desc1"], k1, s["k2"], s["d2"], s["u_right2"],
                                            s["occupied2"], s["cam"], s["mbf"], s["bounds"], s["scale_factors"], th, direction, ori)
        return (np.array([n], np.int32), m.copy())
    return oracle, reference


def case_sbp_map(seed, stereo, th, ratio):
    s = TP.map_scenario(seed, stereo)
    a = (s["proj"], s["view_cos"], s["level1"], s["valid1"], s["obs1"], s["mp_desc1"], s["k2"], s["d2"], s["u_right2"], s["occupied2"],
         s["bounds"], s["scale_factors"], th, ratio)

    def run(fn):
        n, m = fn(*a)
        return (np.array([n], np.int32), _null(m))
    return (lambda: run(O.search_by_projection_map)), (lambda: run(R.search_by_projection_map))


def _bow_keys(angle):
    k = np.zeros(angle.size, O.KP_DTYPE)
    k["angle"], k["x"], k["y"] = angle, 30.0, 30.0
    return k


def case_bow(seed, kf_kf, ratio, ori):
    s = TB.bow_scenario(seed, kf_kf)
    k1, k2 = _bow_keys(s["angle1"]), _bow_keys(s["angle2"])
    bounds = (0.0, 640.0, 0.0, 480.0)

    def oracle():
        n, m12, m21 = O.search_by_bow(s["start1"], s["idx1"], s["start2"], s["idx2"], s["valid1"], s["desc1"], s["angle1"], s["valid2"],
                                      s["desc2"], s["angle2"], 49 if kf_kf else 50, ratio, ori)  # KF-KF: bestDist1 < TH_LOW (:605); KF-F: <= (:228)
        return (np.array([n], np.int32), _null(m12 if kf_kf else m21))

    def reference():
        if kf_kf:
            n, m = R.search_by_bow_keyframes(s["start1"], s["idx1"], s["start2"], s["idx2"], s["valid1"], k1, s["desc1"], s["valid2"], k2,
                                             s["desc2"], bounds, ratio, ori)
        else:
            n, m = R.search_by_bow_frame(s["start1"], s["idx1"], s["start2"], s["idx2"], s["valid1"], k1, s["desc1"], k2, s["desc2"], bounds,
                                         ratio, ori)
        return (np.array([n], np.int32), m.copy())
    return oracle, reference


def case_triangulation(seed, only_stereo, ori):
    s = TT.tri_scenario(seed, only_stereo)
    rng = np.random.default_rng(300 + seed)
    cam = np.array([517.3, 516.5, 318.6, 255.3], F32)
    Cw = np.array([rng.uniform(-1, 1), rng.uniform(-0.5, 0.5), rng.uniform(0.5, 3.0)], F32)
    invz = F32(1.0) / Cw[2]                       # the reference's own order (src/ORBmatcher.cc:668-670)
    ex = F32(F32(cam[0] * Cw[0]) * invz) + cam[2]
    ey = F32(F32(cam[1] * Cw[1]) * invz) + cam[3]
    bounds = (0.0, 640.0, 0.0, 480.0)

    def oracle():
        n, m = O.search_for_triangulation(s["start1"], s["idx1"], s["start2"], s["idx2"], s["kps1"], s["desc1"], s["valid1"], s["stereo1"],
                                          s["kps2"], s["desc2"], s["valid2"], s["stereo2"], s["F12"], float(ex), float(ey),
                                          s["scale_factors2"], s["level_sigma2_2"], ori)
        return (np.array([n], np.int32), m.copy())

    def reference():
        n, m = R.search_for_triangulation(s["start1"], s["idx1"], s["start2"], s["idx2"], s["kps1"], s["desc1"], s["valid1"], s["stereo1"],
                                          s["kps2"], s["desc2"], s["valid2"], s["stereo2"], s["F12"], Cw, cam, bounds, s["scale_factors2"],
                                          only_stereo, ori)
        return (np.array([n], np.int32), m.copy())
    return oracle, reference


def _project(s, rng, th):
    """Map points in the keyframe's camera frame with the gates of src/ORBmatcher.cc:318-362 / :849-888 evaluated the way the reference
    evaluates them (float32 products in the written order, double norm and dot)."""
    x3 = s["x3Dc"].astype(F32)
    n1 = x3.shape[0]
    cam = np.array(s["cam"], F32)
    w, h = s["bounds"][1], s["bounds"][3]
    z = x3[:, 2]
    with np.errstate(divide="ignore", invalid="ignore"):
        invz = (F32(1.0) / z).astype(F32)
        u = (cam[0] * (x3[:, 0] * invz).astype(F32)).astype(F32) + cam[2]
        v = (cam[1] * (x3[:, 1] * invz).astype(F32)).astype(F32) + cam[3]
    dist = np.sqrt((x3.astype(np.float64) ** 2).sum(1)).astype(F32)   # cv::norm: double accumulator, stored in a float
    lo = (dist * rng.uniform(0.3, 1.1, n1)).astype(F32)                # mfMinDistance / mfMaxDistance around the true distance
    hi = (dist * rng.uniform(0.95, 2.5, n1)).astype(F32)
    minmax = np.stack([lo, hi], 1).astype(F32)
    normal = (x3 / np.maximum(dist, 1e-6)[:, None]).astype(F32)
    flip = rng.random(n1) < 0.1
    normal[flip] *= F32(-1.0)                                          # looking from behind: PO . Pn < 0.5 dist
    bad = (rng.random(n1) < 0.05).astype(np.uint8)
    dot = (x3.astype(np.float64) * normal.astype(np.float64)).sum(1)
    ok = (bad == 0) & (z >= 0) & (u >= 0) & (u < w) & (v >= 0) & (v < h)
    ok &= ~((dist < F32(0.8) * lo) | (dist > F32(1.2) * hi))
    ok &= ~(dot < 0.5 * dist.astype(np.float64))
    level = s["level1"].astype(np.int32)
    uvr = np.stack([u, v, (F32(th) * s["scale_factors"][level]).astype(F32)], 1).astype(F32)
    return x3, normal, minmax, level, bad, ok, uvr, invz, cam


def case_sbp_sim3(seed, th):
    s = TP.map_scenario(seed, False)
    s.update({k: v for k, v in TP.scenario(seed, False).items() if k in ("x3Dc", "cam", "mbf")})
    rng = np.random.default_rng(500 + seed)
    x3, normal, minmax, level, bad, ok, uvr, _, cam = _project(s, rng, th)
    found = (rng.random(level.size) < 0.08).astype(np.uint8)
    occ = (rng.random(s["k2"].size) < 0.15).astype(np.uint8)
    valid = (ok & (found == 0)).astype(np.uint8)

    def oracle():
        n, m = O.search_by_projection_sim3(uvr, level, valid, s["mp_desc1"], s["k2"], s["d2"], occ, s["bounds"], 50)
        return (np.array([n], np.int32), m.copy())

    def reference():
        n, m = R.search_by_projection_sim3(x3, normal, minmax, level, bad, found, s["mp_desc1"], s["k2"], s["d2"], occ, cam, s["bounds"],
                                           s["scale_factors"], th)
        return (np.array([n], np.int32), m.copy())
    return oracle, reference


def fuse_surgery(idx, dist, obs1, kf_point_obs, u_right2, n1, entry_point=None, bad_points=(), kf_bad=()):
    """The map surgery that follows Fuse's candidate loop (src/ORBmatcher.cc:947-970) replayed on (bestIdx, bestDist) per list
    entry: the op log the reference harness records, with the state changes of MapPoint::Replace / AddObservation and
    KeyFrame::AddMapPoint (src/MapPoint.cc:83-95, :179-220; src/KeyFrame.cc:209-213).  entry_point[e] = the map point (0 .. n1-1, or
    None for a null pointer) listed at entry e — a point may be listed twice; the keyframe's own points are numbered
    n1 + feature index.  `isBad() || IsInKeyFrame(pKF)` (:849) is evaluated when the entry's turn comes."""
    if entry_point is None:
        entry_point = list(range(n1))
    slot = {i: n1 + i for i in range(len(kf_point_obs)) if kf_point_obs[i] >= 0}   # feature -> point id held by the keyframe
    nobs = {i: int(obs1[i]) for i in range(n1)}
    nobs.update({n1 + i: int(kf_point_obs[i]) for i in slot})
    in_kf = {pid: i for i, pid in slot.items()}                                     # point id -> feature index it observes
    bad = set(int(b) for b in bad_points) | set(n1 + int(j) for j in kf_bad)
    ops, fused = [], 0
    for e, i in enumerate(entry_point):
        if i is None or i in bad or i in in_kf:
            continue
        if idx[e] < 0 or dist[e] > 50:
            continue
        b = int(idx[e])
        held = slot.get(b)
        if held is not None:
            if held not in bad:
                if nobs[held] > nobs[i]:
                    ops.append((2, i, held))          # pMP->Replace(pMPinKF): pMP has no observation to move
                    bad.add(i)
                else:
                    ops.append((2, held, i))          # pMPinKF->Replace(pMP): its observation of the keyframe moves to pMP
                    bad.add(held)
                    f = in_kf.pop(held)
                    slot[f] = i
                    in_kf[i] = f
                    nobs[i] += 2 if u_right2 is not None and u_right2[f] >= 0 else 1
        else:
            ops += [(0, i, b), (1, i, b)]
            in_kf[i] = b
            slot[b] = i
            nobs[i] += 2 if u_right2 is not None and u_right2[b] >= 0 else 1
        fused += 1
    return fused, np.array(ops, np.int64).reshape(-1, 3), slot


def case_fuse(seed, stereo, th):
    s = TP.map_scenario(seed, stereo)
    s.update({k: v for k, v in TP.scenario(seed, stereo).items() if k in ("x3Dc", "cam", "mbf")})
    rng = np.random.default_rng(700 + seed)
    x3, normal, minmax, level, bad, ok, uvr, invz, cam = _project(s, rng, th)
    n1, n2 = level.size, s["k2"].size
    mbf = F32(s["mbf"])
    ur1 = (uvr[:, 0] - (mbf * invz).astype(F32)).astype(F32)
    obs1 = rng.integers(0, 6, n1).astype(np.int32)
    kf_point_obs = np.where(rng.random(n2) < 0.4, rng.integers(1, 6, n2), -1).astype(np.int32)
    inv = (F32(1.0) / (s["scale_factors"] * s["scale_factors"])).astype(F32)
    valid = ok.astype(np.uint8)

    def oracle():
        idx, dist = O.fuse_match(uvr, ur1, level, valid, s["mp_desc1"], s["k2"], s["d2"], s["u_right2"], inv, s["bounds"])
        fused, ops, _ = fuse_surgery(idx, dist, obs1, kf_point_obs, s["u_right2"], n1)
        return (np.array([fused], np.int32), ops)

    def reference():
        n, ops = R.fuse(x3, normal, minmax, level, bad, obs1, s["mp_desc1"], s["k2"], s["d2"], s["u_right2"], kf_point_obs, cam, float(mbf),
                        s["bounds"], s["scale_factors"], th)
        return (np.array([n], np.int32), ops)
    return oracle, reference


def case_stereo(kind, k, nfeat, bf, fx):
    left, right = synth.stereo_pair(kind, k)
    exL, exR = O.Extractor(nfeat, 1.2, 8, 20, 7), O.Extractor(nfeat, 1.2, 8, 20, 7)
    kL, dL = (a.copy() for a in exL.extract(left))
    kR, dR = (a.copy() for a in exR.extract(right))
    h, w = left.shape
    cam = (fx, fx, w / 2.0, h / 2.0)

    def oracle():
        ur, depth = O.stereo_matches(exL, exR, kL, dL, kR, dR, bf, fx)
        return (ur.view(np.uint32).copy(), depth.view(np.uint32).copy())

    def reference():
        lv_l = [exL.level_padded(l)[19:-19, 19:-19] for l in range(8)]
        lv_r = [exR.level_padded(l)[19:-19, 19:-19] for l in range(8)]
        ur, depth = R.stereo_matches(kL, dL, kR, dR, lv_l, lv_r, (w, h), cam, bf, exL.tables()["scale"])
        return (ur.view(np.uint32).copy(), depth.view(np.uint32).copy())
    return oracle, reference


def case_sbp_reloc(seed, th, orb_dist, ori):
    """Relocalisation SearchByProjection (src/ORBmatcher.cc:1472-1599): acceptance distance ORBdist, no depth-sign test, levels
    level-1 .. level+1, sAlreadyFound and bad points skipped, current keypoints that hold ANY map point are not candidates."""
    s = TP.scenario(seed, False)
    rng = np.random.default_rng(800 + seed)
    x3 = s["x3Dc"].astype(F32)
    n1 = x3.shape[0]
    dist = np.sqrt((x3.astype(np.float64) ** 2).sum(1)).astype(F32)
    lo, hi = (dist * rng.uniform(0.3, 1.1, n1)).astype(F32), (dist * rng.uniform(0.95, 2.5, n1)).astype(F32)
    bad, found = (rng.random(n1) < 0.05).astype(np.uint8), (rng.random(n1) < 0.08).astype(np.uint8)
    has = s["valid1"]
    level = np.clip(s["octave1"] + rng.integers(-1, 2, n1), 0, 7).astype(np.int32)
    ok = (has != 0) & (bad == 0) & (found == 0) & ~((dist < F32(0.8) * lo) | (dist > F32(1.2) * hi))
    k1 = np.zeros(n1, O.KP_DTYPE)
    k1["angle"], k1["octave"], k1["x"], k1["y"] = s["angle1"], s["octave1"], 40.0, 40.0

    def oracle():
        n, m = O.search_by_projection_frame(x3, ok.astype(np.uint8), np.ones(n1, np.uint8), s["mp_desc1"], level, s["angle1"], s["k2"], s["d2"],
                                            None, s["occupied2"], s["cam"], s["mbf"], s["bounds"], s["scale_factors"], th, 0, ori, orb_dist, False)
        return (np.array([n], np.int32), _null(m))

    def reference():
        n, m = R.search_by_projection_reloc(x3, has, bad, found, np.stack([lo, hi], 1), level, s["mp_desc1"], k1, s["k2"], s["d2"],
                                            s["occupied2"], s["cam"], s["bounds"], s["scale_factors"], th, orb_dist, ori)
        return (np.array([n], np.int32), m.copy())
    return oracle, reference


def fuse_sim3_replay(idx, dist, entry_point, usable, kf_point_obs, kf_bad, n1):
    """Fuse(pKF, Scw, ...) after its candidate loop (src/ORBmatcher.cc:1077-1095): no Replace; a keypoint that holds a good point is
    reported through vpReplacePoint, a free one takes the point.  usable[e]: the entry passed `!isBad() && !spAlreadyFound.count()`
    and the projection tests."""
    slot = {j: n1 + j for j in range(len(kf_point_obs)) if kf_point_obs[j] >= 0}
    bad = set(n1 + j for j in range(len(kf_point_obs)) if kf_point_obs[j] >= 0 and kf_bad[j])
    ops, rep, fused = [], np.full(len(entry_point), -1, np.int32), 0
    for e, p in enumerate(entry_point):
        if not usable[e] or idx[e] < 0 or dist[e] > 50:
            continue
        b = int(idx[e])
        held = slot.get(b)
        if held is not None:
            if held not in bad:
                rep[e] = held
        else:
            ops += [(0, int(p), b), (1, int(p), b)]
            slot[b] = int(p)
        fused += 1
    return fused, np.array(ops, np.int64).reshape(-1, 3), rep


def case_fuse_sim3(seed, th, scale):
    s = TP.map_scenario(seed, False)
    s.update({k: v for k, v in TP.scenario(seed, False).items() if k in ("x3Dc", "cam", "mbf")})
    rng = np.random.default_rng(900 + seed)
    x3, normal, minmax, level, bad, ok, uvr, invz, cam = _project(s, rng, th)
    n1, n2 = level.size, s["k2"].size
    # Scw = scale * [I | 0]: Rcw = I, tcw = 0, Ow = -0: the same projection; world positions stay x3
    kf_point_obs = np.where(rng.random(n2) < 0.4, rng.integers(1, 6, n2), -1).astype(np.int32)
    kf_bad = ((rng.random(n2) < 0.1) & (kf_point_obs >= 0)).astype(np.uint8)
    entries = np.concatenate([np.arange(n1), rng.integers(0, n1, 40)]).astype(np.int32)  # some points are listed twice
    inv = (F32(1.0) / (s["scale_factors"] * s["scale_factors"])).astype(F32)
    usable = ok[entries]

    def oracle():
        idx, dist = O.fuse_match(uvr[entries], None, level[entries], usable.astype(np.uint8), s["mp_desc1"][entries], s["k2"], s["d2"], None, inv,
                                 s["bounds"])
        fused, ops, rep = fuse_sim3_replay(idx, dist, entries, usable, kf_point_obs, kf_bad, n1)
        return (np.array([fused], np.int32), ops, rep)

    def reference():
        n, ops, rep = R.fuse_sim3(x3, normal, minmax, level, bad, s["mp_desc1"], entries, s["k2"], s["d2"], kf_point_obs, kf_bad, cam,
                                  s["bounds"], s["scale_factors"], scale, th)
        return (np.array([n], np.int32), ops, rep.copy())
    return oracle, reference


def case_sim3(k, th):
    """SearchBySim3 (src/ORBmatcher.cc:1102-1326): both keyframes at the identity pose, points at depth 5 in their own camera, Sim3 =
    the translation that the image shift (7, -4) of synth.shifted_frame corresponds to; expected = two candidate passes
    (O.fuse_match without the gate, TH_HIGH) + the agreement check, with the reference's float order."""
    k1, d1, (h, w) = _extract("tum", k, 1000)
    k2, d2, _ = _extract("tum", k, 1000, shifted=True)
    n1, n2 = k1.size, k2.size
    cam = np.array([500.0, 500.0, 0.5 * w, 0.5 * h], F32)
    sf = O.Extractor(1000, 1.2, 8, 20, 7).tables()["scale"]
    bounds = (0.0, float(w), 0.0, float(h))
    t12 = np.array([-0.07, 0.04, 0.0], F32)
    i1, i2 = np.arange(n1), np.arange(n2)
    has1, has2 = (i1 % 6 != 0).astype(np.uint8), (i2 % 7 != 0).astype(np.uint8)
    bad1, bad2 = (i1 % 17 == 0).astype(np.uint8), (i2 % 17 == 0).astype(np.uint8)

    def world(kp):
        return np.stack([((kp["x"] - cam[2]) * F32(5.0) / cam[0]).astype(F32), ((kp["y"] - cam[3]) * F32(5.0) / cam[1]).astype(F32),
                         np.full(kp.size, 5.0, F32)], 1).astype(F32)
    x1, x2 = world(k1), world(k2)
    matched12 = np.full(n1, -1, np.int32)
    for i in range(4, n1, 25):
        if has1[i]:
            matched12[i] = 1  # keyframe-2 point 1 (has2[1] != 0): already matched, both ends are skipped
    assert has2[1]
    mm = np.array([3.5 / 0.8, 15.5 / 1.2], F32)  # GetMin/MaxDistanceInvariance = 0.8 * min, 1.2 * max

    def one_pass(xsrc, t, has, bad, already, lvl, desc, kdst, ddst):
        # p' = I * p + t with cv::Mat arithmetic: the product rounds to float, then a float addition
        pc = (xsrc + t[None, :]).astype(F32)
        invz = (1.0 / pc[:, 2].astype(np.float64)).astype(F32)
        u = (cam[0] * (pc[:, 0] * invz).astype(F32)).astype(F32) + cam[2]
        v = (cam[1] * (pc[:, 1] * invz).astype(F32)).astype(F32) + cam[3]
        dist = np.sqrt((pc.astype(np.float64) ** 2).sum(1)).astype(F32)
        ok = (has != 0) & ~already & (bad == 0) & (pc[:, 2] >= 0) & (u >= 0) & (u < w) & (v >= 0) & (v < h)
        ok &= ~((dist < F32(0.8) * mm[0]) | (dist > F32(1.2) * mm[1]))
        uvr = np.stack([u, v, (F32(th) * sf[lvl]).astype(F32)], 1).astype(F32)
        inv = (F32(1.0) / (sf * sf)).astype(F32)
        idx, dd = O.fuse_match(uvr, None, lvl, ok.astype(np.uint8), desc, kdst, ddst, None, inv, bounds)
        return np.where((idx >= 0) & (dd <= 100), idx, -1)

    def oracle():
        a1 = matched12 >= 0
        a2 = np.zeros(n2, bool)
        a2[matched12[a1]] = True
        t21 = (-(t12.astype(np.float64))).astype(F32)  # -sR21 * t12 with sR21 = I
        m1 = one_pass(x1, t21, has1, bad1, a1, k1["octave"].astype(np.int32), d1, k2, d2)
        m2 = one_pass(x2, t12, has2, bad2, a2, k2["octave"].astype(np.int32), d2, k1, d1)
        out = matched12.copy()
        found = 0
        for i in range(n1):
            j = m1[i]
            if j >= 0 and m2[j] == i:
                out[i] = j
                found += 1
        return (np.array([found], np.int32), out)

    def reference():
        n, out = R.search_by_sim3(k1, d1, has1, bad1, x1, k1["octave"], k2, d2, has2, bad2, x2, k2["octave"], mm, matched12, cam, bounds, sf,
                                  1.0, t12, th)
        return (np.array([n], np.int32), out.copy())
    return oracle, reference


CASES = {
    "descriptor_distance": lambda: case_descriptor_distance(),
    "features_in_area": lambda: case_features_in_area(),
    "sfi_tum": lambda: case_sfi("tum", 11, 1000, 100, 0.9, True),
    "sfi_kitti_noori": lambda: case_sfi("kitti", 12, 2000, 60, 0.8, False),
    "sfi_dense": lambda: case_sfi("tum", 13, 1000, 100, 0.95, True, random_desc=True),
    "sbp_frame_mono": lambda: case_sbp_frame(1, False, 15.0, 0, True),
    "sbp_frame_fwd": lambda: case_sbp_frame(3, True, 7.0, 1, True),
    "sbp_frame_bwd": lambda: case_sbp_frame(4, True, 7.0, 2, False),
    "sbp_map": lambda: case_sbp_map(1, False, 1.0, 0.8),
    "sbp_map_stereo": lambda: case_sbp_map(2, True, 3.0, 0.8),
    "bow_kf_frame": lambda: case_bow(1, False, 0.7, True),
    "bow_kf_kf": lambda: case_bow(3, True, 0.75, True),
    "bow_kf_kf_noori": lambda: case_bow(4, True, 0.9, False),
    "triangulation": lambda: case_triangulation(1, False, True),
    "triangulation_stereo": lambda: case_triangulation(3, True, True),
    "sbp_sim3": lambda: case_sbp_sim3(11, 10),
    "sbp_sim3_tight": lambda: case_sbp_sim3(12, 4),
    "fuse_stereo": lambda: case_fuse(31, True, 3.0),
    "fuse_mono": lambda: case_fuse(32, False, 4.0),
    "sbp_reloc": lambda: case_sbp_reloc(7, 10.0, 100, True),
    "sbp_reloc_tight": lambda: case_sbp_reloc(8, 3.0, 64, False),
    "fuse_sim3": lambda: case_fuse_sim3(33, 4.0, 2.0),
    "fuse_sim3_unit": lambda: case_fuse_sim3(34, 10.0, 1.0),
    "search_by_sim3": lambda: case_sim3(46, 7.5),
    "stereo_euroc": lambda: case_stereo("euroc", 9, 1200, 47.90639384423901, 435.2046959714599),
}
