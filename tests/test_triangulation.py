"""ORBmatcher::SearchForTriangulation — reference src/ORBmatcher.cc:657-823 with CheckDistEpipolarLine :140-157 (SURVEY.md §8f rank 3;
LocalMapping::CreateNewMapPoints).  CPU: known answers for the oracle (an equal distance later in the node replaces the best,
TH_LOW inclusive, epipole exclusion only for mono-mono pairs, degenerate epipolar line, vbMatched2 never set).  GPU: through the C ABI
against the oracle on real extractions with a fundamental matrix that fits the image shift, bit-exact."""
import numpy as np
import pytest

from oracle import oracle_lib as O
from test_bow import common_nodes, feature_vectors

SF = (np.float32(1.2) ** np.arange(8, dtype=np.float32)).astype(np.float32)
SIGMA2 = (SF * SF).astype(np.float32)


def shift_F(sy):
    """Image 2 = image 1 moved by (sx, sy): the epipolar line of (x1, y1) is y = y1 + sy, i.e. x1' F = [0, 1, -(y1 + sy)]."""
    return np.array([[0, 0, 0], [0, 0, -1], [0, 1, -sy]], np.float32)


def kp(x, y, octave=0, angle=0.0):
    return (x, y, 31.0, angle, 50.0, octave, -1)


def test_oracle_kat():
    k1 = np.array([kp(100, 100)], O.KP_DTYPE)
    k2 = np.array([kp(150, 100), kp(160, 100.5), kp(170, 100), kp(180, 130)], O.KP_DTYPE)
    d1 = np.zeros((1, 32), np.uint8)
    d2 = np.zeros((4, 32), np.uint8)
    d2[0, 0] = 0x03      # 2 bits
    d2[1, 0] = 0x0C      # 2 bits too, later in the node: replaces the best (`dist > bestDist` lets equal through, :734)
    d2[2, 0] = 0xFF      # 8 bits
    d2[3, 0] = 0x00      # 0 bits but 30 px off the epipolar line
    one1, one4, z1, z4 = np.ones(1, np.uint8), np.ones(4, np.uint8), np.zeros(1, np.uint8), np.zeros(4, np.uint8)
    args = ([0, 1], [0], [0, 4], [0, 1, 2, 3], k1, d1, one1, z1, k2, d2, one4, z4, shift_F(0), 1e6, 1e6, SF, SIGMA2)
    n, m12 = O.search_for_triangulation(*args, check_ori=False)
    assert n == 1 and m12[0] == 1
    # the epipole next to keypoint 1 excludes it for a mono-mono pair (:739-745) but not when either side is stereo
    a = list(args)
    a[13], a[14] = 161.0, 100.0
    assert O.search_for_triangulation(*a, check_ori=False)[1][0] == 0
    a[7] = one1
    assert O.search_for_triangulation(*a, check_ori=False)[1][0] == 1
    # TH_LOW is inclusive: 50 bits match, 51 do not
    far = np.zeros((1, 32), np.uint8)
    far[0, :6] = 0xFF
    far[0, 6] = 0x03
    assert O.search_for_triangulation([0, 1], [0], [0, 1], [0], k1, d1, one1, z1, k2[:1], far, one1, z1, shift_F(0), 1e6, 1e6, SF, SIGMA2,
                                      False)[0] == 1
    far[0, 6] = 0x07
    assert O.search_for_triangulation([0, 1], [0], [0, 1], [0], k1, d1, one1, z1, k2[:1], far, one1, z1, shift_F(0), 1e6, 1e6, SF, SIGMA2,
                                      False)[0] == 0
    # a zero fundamental matrix gives den == 0: no pair passes (:151-152)
    assert O.search_for_triangulation(*args[:12], np.zeros((3, 3), np.float32), 1e6, 1e6, SF, SIGMA2, False)[0] == 0
    # vbMatched2 is never set: two first-keyframe features may take the same second-keyframe feature
    k1b = np.array([kp(100, 100), kp(120, 100)], O.KP_DTYPE)
    n, m12 = O.search_for_triangulation([0, 2], [0, 1], [0, 1], [0], k1b, np.zeros((2, 32), np.uint8), [1, 1], [0, 0], k2[:1], d2[:1], [1],
                                        [0], shift_F(0), 1e6, 1e6, SF, SIGMA2, False)
    assert n == 2 and list(m12) == [0, 0]
    # invalid features on either side
    assert O.search_for_triangulation(*args[:6], z1, *args[7:], check_ori=False)[0] == 0


def tri_scenario(seed, only_stereo):
    from orb_slam_cuda_b200 import synth
    rng = np.random.default_rng(seed)
    img = synth.frame("tum", seed)
    sx, sy = int(rng.integers(3, 9)), int(rng.integers(-3, 4))
    cur = np.roll(np.roll(img, sy, axis=0), sx, axis=1)
    ox = O.Extractor(1000, 1.2, 8, 20, 7)
    k1, d1 = (a.copy() for a in ox.extract(img))
    k2, d2 = (a.copy() for a in ox.extract(cur))
    s1, i1, s2, i2 = common_nodes(feature_vectors(d1, 4), feature_vectors(d2, 4))
    st1, st2 = (rng.random(k1.size) < 0.4).astype(np.uint8), (rng.random(k2.size) < 0.4).astype(np.uint8)
    free1, free2 = rng.random(k1.size) < 0.7, rng.random(k2.size) < 0.7
    v1 = (free1 & ((st1 == 1) | (not only_stereo))).astype(np.uint8)
    v2 = (free2 & ((st2 == 1) | (not only_stereo))).astype(np.uint8)
    t = ox.tables()
    return dict(start1=s1, idx1=i1, start2=s2, idx2=i2, kps1=k1, desc1=d1, valid1=v1, stereo1=st1, kps2=k2, desc2=d2, valid2=v2,
                stereo2=st2, F12=shift_F(sy), ex=float(rng.uniform(100, 500)), ey=float(rng.uniform(100, 400)),
                scale_factors2=t["scale"], level_sigma2_2=(t["scale"] * t["scale"]).astype(np.float32))


@pytest.mark.gpu
@pytest.mark.parametrize("seed,only_stereo,ori", [(1, False, True), (2, False, False), (3, True, True), (4, False, True)])
def test_search_for_triangulation_gpu(seed, only_stereo, ori):
    import orb_slam_cuda_b200 as X
    s = tri_scenario(seed, only_stereo)
    n_ref, m_ref = O.search_for_triangulation(s["start1"], s["idx1"], s["start2"], s["idx2"], s["kps1"], s["desc1"], s["valid1"], s["stereo1"],
                                              s["kps2"], s["desc2"], s["valid2"], s["stereo2"], s["F12"], s["ex"], s["ey"],
                                              s["scale_factors2"], s["level_sigma2_2"], ori)
    n, m12 = X.ORBmatcher(0.6, ori).SearchForTriangulation(**s)
    assert n_ref > (20 if only_stereo else 60) and n == n_ref and np.array_equal(m12, m_ref)


@pytest.mark.gpu
def test_search_for_triangulation_edges():
    import orb_slam_cuda_b200 as X
    s = tri_scenario(5, False)
    m = X.ORBmatcher(0.6, True)
    e = np.zeros(0, np.uint32)
    n, m12 = m.SearchForTriangulation(**{**s, "start1": np.zeros(1, np.int32), "idx1": e, "start2": np.zeros(1, np.int32), "idx2": e})
    assert n == 0 and (m12 == -1).all()
    n, m12 = m.SearchForTriangulation(**{**s, "F12": np.zeros((3, 3), np.float32)})
    assert n == 0 and (m12 == -1).all()
    bad = s["kps2"].copy()
    bad["octave"][0] = 9
    with pytest.raises(X.OrbxError):
        m.SearchForTriangulation(**{**s, "kps2": bad})
