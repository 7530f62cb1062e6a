"""The C++ class surface (include/ORBextractor.h, include/ORBmatcher.h) driven the way the reference's Frame.cc /
Tracking.cc drive it, through tests/cpp/frame_harness.cpp.  CPU part: it compiles, links and refuses to run without
a device.  GPU part: its outputs equal the CPU oracle's bit for bit."""
import os
import struct
import subprocess

import numpy as np
import pytest

from oracle import oracle_lib as O
from orb_slam_cuda_b200 import build as B
from orb_slam_cuda_b200 import synth
import orb_slam_cuda_b200 as X

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HARNESS = os.path.join(ROOT, "tests", "cpp", "frame_harness")


@pytest.fixture(scope="module")
def harness():
    B.build()
    B.build_host()
    assert os.path.exists(HARNESS)
    return HARNESS


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def test_class_layer_builds_and_fails_loudly_without_a_device(harness, tmp_path):
    assert os.path.exists(B.CLASSES_SO)
    if _has_gpu():
        pytest.skip("a device is present; the refusal path is checked on CPU boxes")
    raw = tmp_path / "in.raw"
    raw.write_bytes(bytes(2 * 64 * 48))
    r = subprocess.run([harness, "mono", str(raw), "64", "48", "100", str(tmp_path / "o.bin")], capture_output=True, text=True)
    assert r.returncode == 1 and "no CPU path" in r.stderr


class Reader:
    def __init__(self, b):
        self.b, self.o = b, 0

    def i32(self):
        v = struct.unpack_from("<i", self.b, self.o)[0]
        self.o += 4
        return v

    def arr(self, dtype, n):
        a = np.frombuffer(self.b, dtype=dtype, count=n, offset=self.o).copy()
        self.o += a.nbytes
        return a

    def keys(self):
        n = self.i32()
        return self.arr(X.KP_DTYPE, n), self.arr(np.uint8, n * 32).reshape(n, 32)


@pytest.mark.gpu
def test_mono_frame_and_initialization(harness, tmp_path):
    kind, nf = "tum", 1000
    a, b = synth.frame(kind, 40), synth.shifted_frame(kind, 40)
    h, w = a.shape
    raw = tmp_path / "in.raw"
    raw.write_bytes(a.tobytes() + b.tobytes())
    out = tmp_path / "out.bin"
    r = subprocess.run([harness, "mono", str(raw), str(w), str(h), str(nf), str(out)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    rd = Reader(out.read_bytes())
    ox = O.Extractor(nf, 1.2, 8, 20, 7)
    k1, d1 = ox.extract(a)
    lw, lh = rd.i32(), rd.i32()
    assert (lw, lh) == ox.level_size(1)
    lvl = rd.arr(np.uint8, lw * lh).reshape(lh, lw)
    pad = ox.level_padded(1)
    assert np.array_equal(lvl, pad[19:-19, 19:-19])
    assert rd.arr(np.uint8, 1)[0] == pad[18, 18]
    k2, d2 = ox.extract(b)
    gk1, gd1 = rd.keys()
    gk2, gd2 = rd.keys()
    assert np.array_equal(gk1.view(np.uint8), k1.view(np.uint8)) and np.array_equal(gd1, d1)
    assert np.array_equal(gk2.view(np.uint8), k2.view(np.uint8)) and np.array_equal(gd2, d2)
    prev = np.stack([k1["x"], k1["y"]], 1).astype(np.float32)
    n_ref, m_ref, prev_ref = O.search_for_initialization(k1, d1, k2, d2, (0.0, float(w), 0.0, float(h)), prev, 100, 0.9, True)
    assert rd.i32() == n_ref and n_ref > 20
    assert np.array_equal(rd.arr(np.int32, len(k1)), m_ref)
    assert np.array_equal(rd.arr(np.float32, 2 * len(k1)).view(np.uint32), prev_ref.ravel().view(np.uint32))
    t = ox.tables()
    assert rd.i32() == 8
    assert np.array_equal(rd.arr(np.float32, 8), t["scale"]) and np.array_equal(rd.arr(np.float32, 8), t["inv_sigma2"])
    assert rd.i32() == O.descriptor_distance(d1[0], d2[0])
    assert rd.i32() == 302  # the empty-image call left 3 keypoints and the 2-row descriptor matrix untouched


@pytest.mark.gpu
def test_stereo_frame(harness, tmp_path):
    left, right = synth.stereo_pair("euroc", 9)
    h, w = left.shape
    bf, fx = 47.90639384423901, 435.2046959714599
    raw = tmp_path / "in.raw"
    raw.write_bytes(left.tobytes() + right.tobytes())
    out = tmp_path / "out.bin"
    r = subprocess.run([harness, "stereo", str(raw), str(w), str(h), "1200", repr(bf), repr(fx), str(out)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    rd = Reader(out.read_bytes())
    oL, oR = O.Extractor(1200, 1.2, 8, 20, 7), O.Extractor(1200, 1.2, 8, 20, 7)
    kL, dL = oL.extract(left)
    kR, dR = oR.extract(right)
    gkL, gdL = rd.keys()
    gkR, gdR = rd.keys()
    assert np.array_equal(gkL.view(np.uint8), kL.view(np.uint8)) and np.array_equal(gdL, dL)
    assert np.array_equal(gkR.view(np.uint8), kR.view(np.uint8)) and np.array_equal(gdR, dR)
    u_ref, z_ref = O.stereo_matches(oL, oR, kL, dL, kR, dR, float(np.float32(bf)), float(np.float32(fx)))
    u, z = rd.arr(np.float32, len(kL)), rd.arr(np.float32, len(kL))
    assert (u_ref >= 0).sum() > 100
    assert np.array_equal(u.view(np.uint32), u_ref.view(np.uint32)) and np.array_equal(z.view(np.uint32), z_ref.view(np.uint32))


@pytest.mark.gpu
def test_track_with_motion_model_projection(harness, tmp_path):
    """ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono) through the C++ template of include/ORBmatcher.h, with
    a Frame / MapPoint pair shaped like the reference's (tests/cpp/frame_harness.cpp, mode `track`)."""
    kind, nf, th = "tum", 1000, 15.0
    a, b = synth.frame(kind, 41), synth.shifted_frame(kind, 41)
    h, w = a.shape
    raw = tmp_path / "in.raw"
    raw.write_bytes(a.tobytes() + b.tobytes())
    out = tmp_path / "out.bin"
    r = subprocess.run([harness, "track", str(raw), str(w), str(h), str(nf), repr(th), str(out)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    rd = Reader(out.read_bytes())
    k1, d1 = rd.keys()
    k2, d2 = rd.keys()
    n1, n2 = len(k1), len(k2)
    f32 = np.float32
    fx = fy = f32(500.0)
    cx, cy = f32(0.5) * f32(w), f32(0.5) * f32(h)
    i = np.arange(n1)
    z = (f32(2.0) + (i % 17).astype(f32)).astype(f32)
    X0 = ((k1["x"] - cx) * z / fx).astype(f32)
    X1 = ((k1["y"] - cy) * z / fy).astype(f32)
    R = np.array([0.9998, 0, 0.02, 0, 1, 0, -0.02, 0, 0.9998], f32)
    t = np.array([0.01, -0.005, 0.02], f32)
    x3 = np.stack([((R[3 * r] * X0 + R[3 * r + 1] * X1).astype(f32) + R[3 * r + 2] * z).astype(f32) + t[r] for r in range(3)], axis=1).astype(f32)
    has_mp = i % 5 != 0
    valid1 = (has_mp & (i % 11 != 0)).astype(np.uint8)
    obs1 = (i % 7 != 0).astype(np.uint8)
    sf = O.Extractor(nf, 1.2, 8, 20, 7).tables()["scale"]
    n_ref, m_ref = O.search_by_projection_frame(x3, valid1, obs1, d1, k1["octave"], k1["angle"], k2, d2, None, np.zeros(n2, np.uint8),
                                                (fx, fy, cx, cy), 40.0, (0.0, float(w), 0.0, float(h)), sf, th, 0, True)
    assert n_ref > 50
    assert rd.i32() == n_ref
    got = rd.arr(np.int32, n2)
    expect = np.where(m_ref >= 0, m_ref, -1)  # the frame started without map points: NULL-ed and untouched both read as none
    assert np.array_equal(got, expect)
    # second half of the dump: the local-map matcher (src/ORBmatcher.cc:45-129) on the same two frames
    keep = np.flatnonzero(has_mp)                       # vpMapPoints = every existing map point, in index order
    ii = keep
    proj = np.stack([k1["x"][ii] + f32(7.0), k1["y"][ii] - f32(4.0), k1["x"][ii] + f32(7.0) - f32(3.0)], axis=1).astype(f32)
    view_cos = np.where(ii % 2 == 1, f32(0.9995), f32(0.9)).astype(f32)
    valid = ((ii % 3 != 0) & (ii % 13 != 0)).astype(np.uint8)
    n_ref2, m_ref2 = O.search_by_projection_map(proj, view_cos, k1["octave"][ii], valid, obs1[ii], d1[ii], k2, d2, None,
                                                np.zeros(n2, np.uint8), (0.0, float(w), 0.0, float(h)), sf, 3.0, 0.8)
    assert n_ref2 > 50
    assert rd.i32() == n_ref2
    got2 = rd.arr(np.int32, n2)
    assert np.array_equal(got2, np.where(m_ref2 >= 0, keep[np.maximum(m_ref2, 0)], -1))
    # third part: the relocalisation variant (src/ORBmatcher.cc:1472-1599) with the last frame as the keyframe
    Ow = np.array([-((R[c] * t[0] + R[3 + c] * t[1]).astype(f32) + R[6 + c] * t[2]).astype(f32) for c in range(3)], f32)
    Xw = np.stack([X0, X1, z], axis=1).astype(f32)
    PO = (Xw - Ow[None, :]).astype(f32)
    dist3D = np.sqrt((PO.astype(np.float64) ** 2).sum(axis=1)).astype(f32)
    valid3 = (has_mp & (i % 13 != 0) & (i % 9 != 0) & ~(dist3D < f32(3.5)) & ~(dist3D > f32(15.5))).astype(np.uint8)
    occ3 = (np.arange(n2) % 10 == 0).astype(np.uint8)
    n_ref3, m_ref3 = O.search_by_projection_frame(x3, valid3, np.ones(n1, np.uint8), d1, k1["octave"], k1["angle"], k2, d2, None, occ3,
                                                  (fx, fy, cx, cy), 40.0, (0.0, float(w), 0.0, float(h)), sf, 10.0, 0, True, 100, False)
    assert n_ref3 > 50
    assert rd.i32() == n_ref3
    got3 = rd.arr(np.int32, n2)
    assert np.array_equal(got3, np.where(occ3 == 1, 0, np.where(m_ref3 >= 0, m_ref3, -1)))
    # fourth part: the loop-detection variant (src/ORBmatcher.cc:290-403), Scw = scale 2 x the current pose
    S = (f32(2.0) * R).astype(f32)
    scw = f32(np.sqrt(float(S[0]) ** 2 + float(S[1]) ** 2 + float(S[2]) ** 2))
    Rs, ts = (S / scw).astype(f32), ((f32(2.0) * t).astype(f32) / scw).astype(f32)
    pc = np.stack([((Rs[3 * r] * X0 + Rs[3 * r + 1] * X1).astype(f32) + Rs[3 * r + 2] * z).astype(f32) + ts[r] for r in range(3)], axis=1).astype(f32)
    Ow2 = np.array([-((Rs[c] * ts[0] + Rs[3 + c] * ts[1]).astype(f32) + Rs[6 + c] * ts[2]).astype(f32) for c in range(3)], f32)
    PO2 = (Xw - Ow2[None, :]).astype(f32)
    dist2 = np.sqrt((PO2.astype(np.float64) ** 2).sum(axis=1)).astype(f32)
    invz = (f32(1.0) / pc[:, 2]).astype(f32)
    u2 = (fx * (pc[:, 0] * invz).astype(f32)).astype(f32) + cx
    v2 = (fy * (pc[:, 1] * invz).astype(f32)).astype(f32) + cy
    ok = has_mp & (i % 13 != 0) & ~(pc[:, 2] < 0) & (u2 >= 0) & (u2 < f32(w)) & (v2 >= 0) & (v2 < f32(h))
    ok &= ~(dist2 < f32(3.5)) & ~(dist2 > f32(15.5)) & ~(PO2[:, 2].astype(np.float64) < 0.5 * dist2.astype(np.float64))
    uvr = np.stack([u2, v2, (f32(10.0) * sf[k1["octave"]]).astype(f32)], axis=1).astype(f32)
    n_ref4, m_ref4 = O.search_by_projection_sim3(uvr[keep], k1["octave"][keep], ok[keep].astype(np.uint8), d1[keep], k2, d2, occ3,
                                                 (0.0, float(w), 0.0, float(h)), 50)
    assert n_ref4 > 50
    assert rd.i32() == n_ref4
    got4 = rd.arr(np.int32, n2)
    assert np.array_equal(got4, np.where(occ3 == 1, 0, np.where(m_ref4 >= 0, keep[np.maximum(m_ref4, 0)], -1)))
    # SearchBySim3 (src/ORBmatcher.cc:1102-1326) through the template: both keyframes at the identity pose, points at depth 5 in their
    # own camera, Sim3 = translation (-0.07, 0.04, 0) = the image shift (7, -4) at fx = 500
    inv_any = np.ones(8, f32)
    bounds = (0.0, float(w), 0.0, float(h))

    def sim3_pass(ks, ds, skip, already, tvec, kd, dd):
        n = len(ks)
        j = np.arange(n)
        pw = np.stack([((ks["x"] - cx) * f32(5.0) / fx).astype(f32), ((ks["y"] - cy) * f32(5.0) / fy).astype(f32), np.full(n, 5.0, f32)], 1)
        pc = (pw.astype(np.float64) + tvec.astype(np.float64)[None, :]).astype(f32)
        invz = (1.0 / pc[:, 2].astype(np.float64)).astype(f32)
        u = (fx * (pc[:, 0] * invz).astype(f32)).astype(f32) + cx
        v = (fy * (pc[:, 1] * invz).astype(f32)).astype(f32) + cy
        d3 = np.sqrt((pc.astype(np.float64) ** 2).sum(axis=1)).astype(f32)
        ok = (j % skip != 0) & (j % 17 != 0) & ~already & ~(pc[:, 2] < 0) & (u >= 0) & (u < f32(w)) & (v >= 0) & (v < f32(h))
        ok &= ~(d3 < f32(3.5)) & ~(d3 > f32(15.5))
        uvr = np.stack([u, v, (f32(7.5) * sf[ks["octave"]]).astype(f32)], 1).astype(f32)
        idx, dist = O.fuse_match(uvr, None, ks["octave"], ok.astype(np.uint8), ds, kd, dd, None, inv_any, bounds)
        return np.where((idx >= 0) & (dist <= 100), idx, -1)

    t12 = np.array([-0.07, 0.04, 0.0], f32)
    t21 = (-(t12.astype(np.float64))).astype(f32)
    already1 = (i % 25 == 4) & (i % 6 != 0)
    already2 = np.zeros(n2, bool)
    already2[1] = already1.any()
    vn1 = sim3_pass(k1, d1, 6, already1, t21, k2, d2)
    vn2 = sim3_pass(k2, d2, 7, already2, t12, k1, d1)
    want = np.where(already1, 1, -1)
    agree = (vn1 >= 0) & (vn2[np.maximum(vn1, 0)] == i)
    want = np.where(agree, vn1, want)
    assert agree.sum() > 50
    assert rd.i32() == int(agree.sum())
    assert np.array_equal(rd.arr(np.int32, n1), want)
    # fifth part: SearchByBoW(keyframe, frame) (src/ORBmatcher.cc:159-288), node id = 5 descriptor bits
    from test_bow import common_nodes, feature_vectors
    s1, i1, s2, i2 = common_nodes(feature_vectors(d1, 5), feature_vectors(d2, 5))
    valid_kf = (has_mp & (i % 13 != 0)).astype(np.uint8)
    n_ref5, _, m21_5 = O.search_by_bow(s1, i1, s2, i2, valid_kf, d1, k1["angle"], np.ones(n2, np.uint8), d2, k2["angle"], 50, 0.7, True)
    assert n_ref5 > 50
    assert rd.i32() == n_ref5
    assert np.array_equal(rd.arr(np.int32, n2), np.where(m21_5 >= 0, m21_5, -1))
    # sixth part: SearchByBoW(keyframe, keyframe) (:522-655); keypoint j of the second keyframe holds point j % n1
    j = np.arange(n2)
    valid_kc = ((j % 4 != 0) & ((j % n1) % 13 != 0)).astype(np.uint8)
    n_ref6, m12_6, _ = O.search_by_bow(s1, i1, s2, i2, valid_kf, d1, k1["angle"], valid_kc, d2, k2["angle"], 49, 0.75, True)
    assert n_ref6 > 50
    assert rd.i32() == n_ref6
    assert np.array_equal(rd.arr(np.int32, n1), np.where(m12_6 >= 0, np.maximum(m12_6, 0) % n1, -1))


@pytest.mark.gpu
def test_compute_bow_and_search_by_bow(harness, tmp_path):
    """ORBVocabulary::loadFromTextFile + Frame::ComputeBoW + ORBmatcher::SearchByBoW through include/ORBVocabulary.h and
    include/ORBmatcher.h (tests/cpp/frame_harness.cpp, mode `bow`): BowVector doubles, FeatureVector lists and the matches equal
    the oracle's on the same text vocabulary."""
    from test_bow import common_nodes
    from test_vocabulary import make_vocabulary
    kind, nf = "tum", 1000
    a, b = synth.frame(kind, 41), synth.shifted_frame(kind, 41)
    h, w = a.shape
    raw = tmp_path / "in.raw"
    raw.write_bytes(a.tobytes() + b.tobytes())
    # a vocabulary whose level-1..3 centroids are ORB descriptors of a third image, so real descriptors spread over the tree
    _, seed_desc = O.Extractor(1000, 1.2, 8, 20, 7).extract(synth.frame(kind, 7))
    voc = make_vocabulary(21, 8, 3, p_early=0.05)
    rng = np.random.default_rng(5)
    voc.desc[1:] = seed_desc[rng.integers(0, seed_desc.shape[0], voc.desc.shape[0] - 1)]
    path = tmp_path / "ORBvoc.txt"
    voc.to_text(path)
    ref = O.Vocabulary.from_text(path)
    out = tmp_path / "out.bin"
    r = subprocess.run([harness, "bow", str(raw), str(w), str(h), str(nf), str(path), str(out), "-4"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    rd = Reader(out.read_bytes())
    assert rd.i32() == 0                       # a missing file: loadFromTextFile returns false
    assert rd.i32() == int(ref.is_leaf[1:].sum())
    frames = []
    for _ in range(2):
        k, d = rd.keys()
        ref_bow, ref_fv = ref.transform(d, 4)
        nb = rd.i32()
        assert nb == len(ref_bow) > 20
        for word, val in ref_bow.items():
            assert rd.i32() == word
            assert rd.arr(np.uint64, 1)[0] == np.float64(val).view(np.uint64)
        assert rd.i32() == len(ref_fv)
        for node, idx in ref_fv.items():
            assert rd.i32() == node and rd.i32() == len(idx)
            assert list(rd.arr(np.uint32, len(idx))) == idx
        frames.append((k, d, {n: np.array(v, np.uint32) for n, v in ref_fv.items()}))
    assert rd.i32() == 1                       # transform(vector<cv::Mat>) == transform_rows
    (k1, d1, fv1), (k2, d2, fv2) = frames
    s1, i1, s2, i2 = common_nodes(fv1, fv2)
    i = np.arange(len(k1))
    valid1 = ((i % 5 != 0) & (i % 13 != 0)).astype(np.uint8)
    n_ref, _, m21 = O.search_by_bow(s1, i1, s2, i2, valid1, d1, k1["angle"], np.ones(len(k2), np.uint8), d2, k2["angle"], 50, 0.7, True)
    assert n_ref > 30
    assert rd.i32() == n_ref
    assert np.array_equal(rd.arr(np.int32, len(k2)), np.where(m21 >= 0, m21, -1))
    # SearchForTriangulation(pKF1, pKF2, F12, vMatchedPairs, false) through the template (src/ORBmatcher.cc:657-823); the second image
    # is the first moved by (7, -4), so the harness passes F12 for the lines y = y1 - 4 and a camera 0.5 m to the right
    f32 = np.float32
    i, j = np.arange(len(k1)), np.arange(len(k2))
    v1, t1 = (i % 5 == 0).astype(np.uint8), (i % 3 == 0).astype(np.uint8)
    v2, t2 = (j % 6 != 0).astype(np.uint8), (j % 4 == 0).astype(np.uint8)
    invz = f32(1.0) / f32(1e-3)
    ex = f32(f32(f32(500.0) * f32(-0.5)) * invz) + f32(0.5) * f32(w)
    ey = f32(f32(f32(500.0) * f32(0.01)) * invz) + f32(0.5) * f32(h)
    sf = (np.float32(1.2) ** np.arange(8)).astype(np.float32)
    ox = O.Extractor(nf, 1.2, 8, 20, 7)
    sf, sg = ox.tables()["scale"], ox.tables()["sigma2"]
    F12 = np.array([[0, 0, 0], [0, 0, -1], [0, 1, 4]], np.float32)
    n_tri, m_tri = O.search_for_triangulation(s1, i1, s2, i2, k1, d1, v1, t1, k2, d2, v2, t2, F12, float(ex), float(ey), sf, sg, False)
    assert n_tri > 10
    assert rd.i32() == n_tri
    assert rd.i32() == n_tri
    pairs = rd.arr(np.int32, 2 * n_tri).reshape(-1, 2)
    want = np.flatnonzero(m_tri >= 0)
    assert np.array_equal(pairs[:, 0], want) and np.array_equal(pairs[:, 1], m_tri[want])


@pytest.mark.gpu
def test_fuse_both_overloads_with_map_surgery(harness, tmp_path):
    """ORBmatcher::Fuse(pKF, vpMapPoints, th) and Fuse(pKF, Scw, vpPoints, th, vpReplacePoint) through the templates of
    include/ORBmatcher.h with mock KeyFrame / MapPoint classes that record the map surgery (tests/cpp/frame_harness.cpp, mode
    `fuse`).  Expected: the oracle's candidate loop (O.fuse_match) followed by the replay of src/ORBmatcher.cc:947-970 /
    :1077-1095 (tests/ref_matcher_cases.fuse_surgery — the replay that tests/test_oracle_ref_matcher.py pins to the reference's
    own Fuse compiled from /root/reference).  nFused, the op log in order, the keyframe's final slots and vpReplacePoint equal."""
    import ref_matcher_cases as RC
    kind, nf, th = "tum", 1000, 3.0
    a, b = synth.frame(kind, 43), synth.shifted_frame(kind, 43)
    h, w = a.shape
    raw = tmp_path / "in.raw"
    raw.write_bytes(a.tobytes() + b.tobytes())
    out = tmp_path / "out.bin"
    r = subprocess.run([harness, "fuse", str(raw), str(w), str(h), str(nf), repr(th), str(out)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    rd = Reader(out.read_bytes())
    k1, d1 = rd.keys()
    k2, d2 = rd.keys()
    n1, n2 = len(k1), len(k2)
    f32 = np.float32
    fx = fy = f32(500.0)
    cx, cy = f32(0.5) * f32(w), f32(0.5) * f32(h)
    t = O.Extractor(nf, 1.2, 8, 20, 7).tables()
    sf, inv_sigma2 = t["scale"], t["inv_sigma2"]
    i = np.arange(n1)
    # the harness' world points (float expressions in its order), projected as project_for_fuse does with the identity pose
    X0 = (((k1["x"] + f32(7.0)).astype(f32) - cx).astype(f32) * f32(5.0) / fx).astype(f32)
    X1 = (((k1["y"] - f32(4.0)).astype(f32) - cy).astype(f32) * f32(5.0) / fy).astype(f32)
    z = np.full(n1, 5.0, f32)
    invz = (f32(1.0) / z).astype(f32)
    u = (fx * (X0 * invz).astype(f32)).astype(f32) + cx
    v = (fy * (X1 * invz).astype(f32)).astype(f32) + cy
    dist3 = np.sqrt(X0.astype(np.float64) ** 2 + X1.astype(np.float64) ** 2 + 25.0).astype(f32)
    nz = np.where(i % 19 == 0, -1.0, 1.0)
    ok = (u >= 0) & (u < w) & (v >= 0) & (v < h) & ~((dist3 < f32(3.5)) | (dist3 > f32(15.5))) & ~(5.0 * nz < 0.5 * dist3.astype(np.float64))
    level = k1["octave"].astype(np.int32)
    u_right2 = np.full(n2, -1.0, f32)
    u_right2[::3] = k2["x"][::3] - f32(8.0)
    kf_point_obs = np.where(np.arange(n2) % 4 == 1, 1 + np.arange(n2) % 5, -1).astype(np.int32)
    kf_bad = [j for j in range(n2) if j % 4 == 1 and j % 29 == 1]
    obs1 = (i % 6).astype(np.int32)
    bad_points = [p for p in range(n1) if p % 23 == 0]
    entries = [None if p % 5 == 0 else p for p in range(n1)] + list(range(1, n1, 37))
    bounds = (0.0, float(w), 0.0, float(h))
    assert len(set(e for e in entries if e is not None)) < len([e for e in entries if e is not None])  # some points are listed twice

    def gather(ent, gate):
        e_idx = np.array([0 if e is None else e for e in ent])
        valid = np.array([e is not None for e in ent]) & ok[e_idx]
        uvr = np.stack([u[e_idx], v[e_idx], (f32(th) * sf[level[e_idx]]).astype(f32)], 1).astype(f32)
        ur1 = (u[e_idx] - (f32(40.0) * invz[e_idx]).astype(f32)).astype(f32) if gate else None
        return O.fuse_match(uvr, ur1, level[e_idx], valid.astype(np.uint8), d1[e_idx], k2, d2, u_right2 if gate else None, inv_sigma2, bounds)

    # ---- overload 1 ----
    idx, dist = gather(entries, True)
    fused, ops, slot = RC.fuse_surgery(idx, dist, obs1, kf_point_obs, u_right2, n1, entries, bad_points, kf_bad)
    assert fused > 100 and (ops[:, 0] == 2).sum() > 10 and (ops[:, 0] == 0).sum() > 50
    assert rd.i32() == fused
    n_ops = rd.i32()
    got_ops = rd.arr(np.int32, 3 * n_ops).reshape(-1, 3)
    assert np.array_equal(got_ops, ops)
    got_slots = rd.arr(np.int32, n2)
    want_slots = np.array([slot.get(j, -1) for j in range(n2)], np.int32)
    assert np.array_equal(got_slots, want_slots)
    assert rd.i32() == len(entries)
    assert (rd.arr(np.int32, len(entries)) == -1).all()  # the first overload has no vpReplacePoint
    # ---- overload 2 (Scw = 2 * identity: scale 2, rotation I, translation 0): no chi-square gate, no Replace ----
    entries2 = [0 if e is None else e for e in entries]  # the harness puts (bad) point 0 where overload 1 had null pointers
    idx, dist = gather(entries2, False)
    in_kf = set()
    slot2 = {j: n1 + j for j in range(n2) if kf_point_obs[j] >= 0}
    ops2, replaced, fused2 = [], np.full(len(entries2), -1, np.int32), 0
    for e, p in enumerate(entries2):
        if p in bad_points or not ok[p] or idx[e] < 0 or dist[e] > 50:   # spAlreadyFound holds keyframe points only: never a candidate
            continue
        bI = int(idx[e])
        held = slot2.get(bI)
        if held is not None:
            if held >= n1 and (held - n1) in kf_bad:
                pass
            else:
                replaced[e] = held
        else:
            ops2 += [(0, p, bI), (1, p, bI)]
            slot2[bI] = p
        fused2 += 1
    assert fused2 > 100 and (replaced >= 0).sum() > 10
    assert rd.i32() == fused2
    n_ops = rd.i32()
    assert np.array_equal(rd.arr(np.int32, 3 * n_ops).reshape(-1, 3), np.array(ops2, np.int64).reshape(-1, 3))
    assert np.array_equal(rd.arr(np.int32, n2), np.array([slot2.get(j, -1) for j in range(n2)], np.int32))
    assert rd.i32() == len(entries2)
    assert np.array_equal(rd.arr(np.int32, len(entries2)), replaced)


REF_UNCHANGED = os.path.join(ROOT, "oracle", "_ref", "ref_frame_unchanged")


@pytest.mark.gpu
def test_reference_frame_cc_runs_unchanged_on_the_class_surface(tmp_path):
    """BASELINE.json north_star: "Frame.cc ... and ComputeStereoMatches run unchanged on top of it".  oracle/_ref/ref_frame_unchanged is the
    REFERENCE's own src/Frame.cc (compiled from /root/reference by oracle/ref_matcher/Makefile) linked against this repository's
    ORBextractor / ORBmatcher classes and liborbx.so.  Its stereo constructor extracts both eyes through ORBextractor::operator() on two
    threads and then runs its unchanged ComputeStereoMatches body over mvImagePyramid (host copies of the device pyramid) and
    ORBmatcher::DescriptorDistance (one device call per pair).  Keypoints, descriptors, mvuRight / mvDepth (float bits) must equal the
    oracle's — and therefore orbx_stereo_match, which test_matcher_gpu compares with the same oracle — and the reference's mGrid must
    equal the grid orbx_frame bins on the device."""
    if not os.path.exists(REF_UNCHANGED):
        pytest.skip("oracle/_ref/ref_frame_unchanged is built where /root/reference exists")
    w, h, nf, bf, fx = 752, 480, 1200, 47.90639384423901, 435.2046959714599
    left, right = synth.stereo_pair("euroc", 9)
    raw = tmp_path / "in.raw"
    raw.write_bytes(left.tobytes() + right.tobytes())
    out = tmp_path / "out.bin"
    r = subprocess.run([REF_UNCHANGED, str(raw), str(w), str(h), str(nf), repr(bf), repr(fx), str(out)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr
    rd = Reader(out.read_bytes())
    kL, dL = rd.keys()
    kR, dR = rd.keys()
    exL, exR = O.Extractor(nf, 1.2, 8, 20, 7), O.Extractor(nf, 1.2, 8, 20, 7)
    okL, odL = exL.extract(left)
    okR, odR = exR.extract(right)
    assert np.array_equal(kL.view(np.uint8), okL.view(np.uint8)) and np.array_equal(dL, odL)
    assert np.array_equal(kR.view(np.uint8), okR.view(np.uint8)) and np.array_equal(dR, odR)
    ur, depth = rd.arr(np.float32, len(kL)), rd.arr(np.float32, len(kL))
    our, odepth = O.stereo_matches(exL, exR, okL.copy(), odL.copy(), okR.copy(), odR.copy(), bf, fx)
    assert (ur >= 0).sum() > 300
    assert np.array_equal(ur.view(np.uint32), our.view(np.uint32)) and np.array_equal(depth.view(np.uint32), odepth.view(np.uint32))
    gL, gR = X.ORBextractor(nf, 1.2, 8, 20, 7, w, h), X.ORBextractor(nf, 1.2, 8, 20, 7, w, h)
    k1, d1 = gL(left)
    k2, d2 = gR(right)
    dur, ddepth = X.ComputeStereoMatches(gL, gR, k1, d1, k2, d2, bf, fx)
    assert np.array_equal(dur.view(np.uint32), ur.view(np.uint32)) and np.array_equal(ddepth.view(np.uint32), depth.view(np.uint32))
    # the reference's mGrid (Frame::AssignFeaturesToGrid) against the device-resident frame's grid
    cells = []
    for _ in range(64 * 48):
        n = rd.i32()
        cells.append(rd.arr(np.int32, n))
    cs, so = X.DeviceFrame(kL, dL, (0.0, float(w), 0.0, float(h))).grid()
    assert np.array_equal(np.concatenate([[0], np.cumsum([c.size for c in cells])]).astype(np.int32), cs)
    assert np.array_equal(np.concatenate(cells), so)


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["mono", "stereo", "track", "fuse"])
def test_opencv_branches_of_the_class_layer(harness, tmp_path, mode):
    """include/ORBmatcher.h and host/ORBextractor.cc have `#if ORBX_HAVE_OPENCV` branches — the ones a real ORB-SLAM2 build uses: the
    reference's own cv::Mat expressions for the pose arithmetic instead of the stand-in's scalar code.  tests/cpp/frame_harness_cv is the
    same harness compiled with those branches on (the test-only <opencv2/...> of oracle/ref_matcher/shim as OpenCV: products, dot and
    norm with a double accumulator like OpenCV's CV_32F paths).  Both builds must write the same bytes in every mode."""
    cv = harness + "_cv"
    if not os.path.exists(cv):
        pytest.skip("frame_harness_cv not built")
    kind = "euroc" if mode == "stereo" else "tum"
    if mode == "stereo":
        a, b = synth.stereo_pair(kind, 4)
    else:
        a, b = synth.frame(kind, 45), synth.shifted_frame(kind, 45)
    h, w = a.shape
    raw = tmp_path / "in.raw"
    raw.write_bytes(a.tobytes() + b.tobytes())
    extra = {"mono": [], "stereo": ["47.9", "435.2"], "track": ["15.0"], "fuse": ["3.0"]}[mode]
    outs = []
    for exe in (harness, cv):
        out = tmp_path / (os.path.basename(exe) + ".bin")
        r = subprocess.run([exe, mode, str(raw), str(w), str(h), "1000"] + extra + [str(out)], capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stderr
        outs.append(out.read_bytes())
    assert len(outs[0]) > 1000 and outs[0] == outs[1]
