"""SearchByProjection(CurrentFrame, LastFrame, th, bMono) — reference src/ORBmatcher.cc:1328-1470.
CPU: known-answer cases that pin the oracle's reading of the reference loop.  GPU: the CUDA path through the C ABI against
the oracle on keypoints / descriptors of real extractions, bit-exact (integer assignment + match count)."""
import numpy as np
import pytest

from oracle import oracle_lib as O

BOUNDS = (0.0, 640.0, 0.0, 480.0)
CAM = (500.0, 500.0, 320.0, 240.0)
SF = np.float32(1.2) ** np.arange(8, dtype=np.float32)


def kp(x, y, octave=0, angle=0.0):
    return (x, y, 31.0, angle, 50.0, octave, -1)


def point_at(u, v, z=4.0):
    return ((u - CAM[2]) * z / CAM[0], (v - CAM[3]) * z / CAM[1], z)


def test_oracle_kat_competition_and_observations():
    """Two last-frame points project next to the same current keypoint.  With observations the first one to be processed
    keeps it and the second falls back to its second-best candidate; a point without observations (a temporal stereo point)
    does not block the keypoint, so the later iteration overwrites the assignment and nmatches counts both events."""
    k2 = np.array([kp(100, 100), kp(103, 100), kp(300, 300)], dtype=O.KP_DTYPE)
    d2 = np.zeros((3, 32), np.uint8)
    d2[1, 0] = 0x0F          # 4 bits away from the zero descriptor
    d2[2] = 0xFF
    x3 = np.array([point_at(101, 100), point_at(101.5, 100)], np.float32)
    mp = np.zeros((2, 32), np.uint8)
    one, zero = np.ones(2, np.uint8), np.zeros(2, np.uint8)
    args = dict(octave1=[0, 0], angle1=[0, 0], k2=k2, d2=d2, u_right2=None, occupied2=np.zeros(3, np.uint8), cam=CAM, mbf=40.0,
                bounds=BOUNDS, scale_factors=SF, th=7.0, direction=0, check_ori=False)
    n, m21 = O.search_by_projection_frame(x3, one, one, mp, **args)
    assert n == 2 and list(m21) == [0, 1, -1]
    n, m21 = O.search_by_projection_frame(x3, one, zero, mp, **args)
    assert n == 2 and list(m21) == [1, -1, -1]
    # behind the camera / outside the image / invalid points are skipped
    x3b = np.array([(0, 0, -1.0), point_at(-50, 100)], np.float32)
    n, m21 = O.search_by_projection_frame(x3b, one, one, mp, **args)
    assert n == 0 and list(m21) == [-1, -1, -1]
    n, m21 = O.search_by_projection_frame(x3, zero, one, mp, **args)
    assert n == 0
    # occupied keypoints are never candidates; TH_HIGH = 100 bounds the distance
    occ = np.array([1, 0, 0], np.uint8)
    n, m21 = O.search_by_projection_frame(x3[:1], one[:1], one[:1], mp[:1], **{**args, "occupied2": occ, "octave1": [0], "angle1": [0]})
    assert n == 1 and list(m21) == [-1, 0, -1]
    far = np.full((1, 32), 0xFF, np.uint8)
    n, m21 = O.search_by_projection_frame(x3[:1], one[:1], one[:1], far, **{**args, "octave1": [0], "angle1": [0]})
    assert n == 0


def test_oracle_kat_levels_stereo_and_rotation():
    k2 = np.array([kp(200, 200, 0), kp(201, 200, 2), kp(202, 200, 4)], dtype=O.KP_DTYPE)
    d2 = np.zeros((3, 32), np.uint8)
    d2[0, 0], d2[1, 0], d2[2, 0] = 1, 3, 7
    x3 = np.array([point_at(200, 200)], np.float32)
    one = np.ones(1, np.uint8)
    args = dict(angle1=[0], k2=k2, d2=d2, u_right2=None, occupied2=np.zeros(3, np.uint8), cam=CAM, mbf=40.0, bounds=BOUNDS,
                scale_factors=SF, th=15.0, check_ori=False)
    mp = np.zeros((1, 32), np.uint8)
    # octave 2: +-1 levels -> only keypoint 1; forward (>= 2) -> 1 (closer than 2); backward (0..2) -> 0
    assert list(O.search_by_projection_frame(x3, one, one, mp, octave1=[2], direction=0, **args)[1]) == [-1, 0, -1]
    assert list(O.search_by_projection_frame(x3, one, one, mp, octave1=[2], direction=1, **args)[1]) == [-1, 0, -1]
    assert list(O.search_by_projection_frame(x3, one, one, mp, octave1=[2], direction=2, **args)[1]) == [0, -1, -1]
    # stereo: a right coordinate further than the radius from u - mbf/z rejects the candidate (mvuRight <= 0: no check)
    ur_pred = 200.0 - 40.0 / 4.0
    ur = np.array([ur_pred + 100.0, -1.0, -1.0], np.float32)
    a2 = {**args, "u_right2": ur}
    assert list(O.search_by_projection_frame(x3, one, one, mp, octave1=[1], direction=2, **a2)[1]) == [-1, -1, -1]
    ur[0] = ur_pred + 1.0
    assert list(O.search_by_projection_frame(x3, one, one, mp, octave1=[1], direction=2, **a2)[1]) == [0, -1, -1]
    # rotation histogram: 12 matches rotate by ~0 degrees, one by 90: the odd one is set to NULL (-2) and not counted
    n2 = 13
    k2 = np.array([kp(50 + 30 * i, 100, 0, 90.0 if i == 12 else 0.0) for i in range(n2)], dtype=O.KP_DTYPE)
    x3 = np.array([point_at(50 + 30 * i, 100) for i in range(n2)], np.float32)
    ones = np.ones(n2, np.uint8)
    n, m21 = O.search_by_projection_frame(x3, ones, ones, np.zeros((n2, 32), np.uint8), octave1=[0] * n2, angle1=[0.0] * n2, k2=k2,
                                          d2=np.zeros((n2, 32), np.uint8), u_right2=None, occupied2=np.zeros(n2, np.uint8), cam=CAM,
                                          mbf=40.0, bounds=BOUNDS, scale_factors=SF, th=7.0, direction=0, check_ori=True)
    assert n == 12 and list(m21) == list(range(12)) + [-2]


def scenario(seed, stereo, kind="tum"):
    """Last frame = an extraction; current frame = the same image shifted by a few pixels (so true correspondences exist);
    map points = last-frame keypoints back-projected at random depths and moved into the current camera frame."""
    from orb_slam_cuda_b200 import synth
    rng = np.random.default_rng(seed)
    img = synth.frame(kind, seed)
    sx, sy = int(rng.integers(2, 7)), int(rng.integers(-4, 5))
    cur = np.roll(np.roll(img, sy, axis=0), sx, axis=1)
    ox = O.Extractor(1000, 1.2, 8, 20, 7)
    k1, d1 = ox.extract(img)
    k1, d1 = k1.copy(), d1.copy()
    k2, d2 = ox.extract(cur)
    k2, d2 = k2.copy(), d2.copy()
    h, w = img.shape
    cam = (517.3, 516.5, 318.6, 255.3)
    n1, n2 = k1.size, k2.size
    z = rng.uniform(1.5, 20.0, n1).astype(np.float32)
    z[rng.random(n1) < 0.03] *= -1  # some points behind the camera
    u = k1["x"] + sx + rng.normal(0, 1.5, n1).astype(np.float32)
    v = k1["y"] + sy + rng.normal(0, 1.5, n1).astype(np.float32)
    x3 = np.stack([(u - cam[2]) * z / cam[0], (v - cam[3]) * z / cam[1], z], axis=1).astype(np.float32)
    mp = d1.copy()
    flips = rng.integers(0, 256, size=mp.shape, dtype=np.uint8) & rng.integers(0, 256, size=mp.shape, dtype=np.uint8) & \
        rng.integers(0, 256, size=mp.shape, dtype=np.uint8) & rng.integers(0, 256, size=mp.shape, dtype=np.uint8)
    mp ^= flips  # ~16 flipped bits per descriptor
    valid1 = (rng.random(n1) < 0.8).astype(np.uint8)
    obs1 = (rng.random(n1) < 0.85).astype(np.uint8)
    occupied2 = (rng.random(n2) < 0.1).astype(np.uint8)
    mbf = 40.0
    ur = None
    if stereo:
        ur = np.where(rng.random(n2) < 0.6, k2["x"] - mbf / rng.uniform(1.5, 20.0, n2), -1.0).astype(np.float32)
    return dict(x3Dc=x3, valid1=valid1, obs1=obs1, mp_desc1=mp, octave1=k1["octave"], angle1=k1["angle"], k2=k2, d2=d2,
                u_right2=ur, occupied2=occupied2, cam=cam, mbf=mbf, bounds=(0.0, float(w), 0.0, float(h)),
                scale_factors=ox.tables()["scale"])


@pytest.mark.gpu
@pytest.mark.parametrize("seed,stereo,th,direction,ori", [(1, False, 15.0, 0, True), (2, False, 7.0, 0, True), (3, True, 7.0, 1, True),
                                                          (4, True, 7.0, 2, True), (5, True, 15.0, 0, False), (6, False, 30.0, 0, True)])
def test_search_by_projection_gpu(seed, stereo, th, direction, ori):
    import orb_slam_cuda_b200 as X
    s = scenario(seed, stereo)
    n_ref, m_ref = O.search_by_projection_frame(s["x3Dc"], s["valid1"], s["obs1"], s["mp_desc1"], s["octave1"], s["angle1"], s["k2"],
                                                s["d2"], s["u_right2"], s["occupied2"], s["cam"], s["mbf"], s["bounds"],
                                                s["scale_factors"], th, direction, ori)
    m = X.ORBmatcher(0.9, ori)
    n, m21 = m.SearchByProjection(s["x3Dc"], s["valid1"], s["obs1"], s["mp_desc1"], s["octave1"], s["angle1"], s["k2"], s["d2"],
                                  s["u_right2"], s["occupied2"], s["cam"], s["mbf"], s["bounds"], s["scale_factors"], th, direction)
    assert n_ref > 50, "the scenario must produce matches"
    assert n == n_ref
    assert np.array_equal(m21, m_ref)


def test_oracle_kat_relocalisation_knobs():
    """SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist), src/ORBmatcher.cc:1472-1599: ORBdist replaces TH_HIGH
    and a point behind the camera is not rejected (only the frame-to-frame routine tests invzc < 0)."""
    k2 = np.array([kp(100, 100)], dtype=O.KP_DTYPE)
    d2 = np.zeros((1, 32), np.uint8)
    d2[0, :9] = 0xFF  # 72 bits from the zero descriptor
    args = dict(valid1=[1], obs1=[1], mp_desc1=np.zeros((1, 32), np.uint8), octave1=[0], angle1=[0], k2=k2, d2=d2, u_right2=None,
                occupied2=[0], cam=CAM, mbf=40.0, bounds=BOUNDS, scale_factors=SF, th=10.0, direction=0, check_ori=False)
    front = np.array([point_at(101, 100)], np.float32)
    behind = (-front).astype(np.float32)  # same projection, negative depth
    assert O.search_by_projection_frame(front, **args)[0] == 1
    assert O.search_by_projection_frame(front, orb_dist=64, **args)[0] == 0
    assert O.search_by_projection_frame(behind, **args)[0] == 0
    assert O.search_by_projection_frame(behind, depth_test=False, **args)[0] == 1


@pytest.mark.gpu
@pytest.mark.parametrize("seed,th,orb_dist", [(7, 10.0, 100), (8, 3.0, 64)])
def test_search_by_projection_relocalisation_gpu(seed, th, orb_dist):
    import orb_slam_cuda_b200 as X
    s = scenario(seed, False)
    ones = np.ones_like(s["obs1"])
    a = (s["x3Dc"], s["valid1"], ones, s["mp_desc1"], s["octave1"], s["angle1"], s["k2"], s["d2"], None, s["occupied2"], s["cam"],
         s["mbf"], s["bounds"], s["scale_factors"], th, 0)
    n_ref, m_ref = O.search_by_projection_frame(*a, True, orb_dist, False)
    n, m21 = X.ORBmatcher(0.9, True).SearchByProjection(*a, orb_dist=orb_dist, depth_test=False)
    assert n_ref > 50 and n == n_ref and np.array_equal(m21, m_ref)
    n_dt, _ = O.search_by_projection_frame(*a, True, orb_dist, True)
    assert n_dt != n_ref or (s["x3Dc"][:, 2] > 0).all()  # the points behind the camera make a difference


@pytest.mark.gpu
def test_search_by_projection_empty_and_errors():
    import orb_slam_cuda_b200 as X
    m = X.ORBmatcher(0.9, True)
    k2 = np.array([kp(100, 100)], dtype=X.KP_DTYPE)
    n, m21 = m.SearchByProjection(np.zeros((0, 3), np.float32), [], [], np.zeros((0, 32), np.uint8), [], [], k2,
                                  np.zeros((1, 32), np.uint8), None, [0], CAM, 40.0, BOUNDS, SF, 7.0)
    assert n == 0 and list(m21) == [-1]
    with pytest.raises(Exception):
        m.SearchByProjection(np.zeros((1, 3), np.float32), [1], [1], np.zeros((1, 32), np.uint8), [9], [0.0], k2,
                             np.zeros((1, 32), np.uint8), None, [0], CAM, 40.0, BOUNDS, SF, 7.0)


# ---- SearchByProjection(F, vpMapPoints, th) — src/ORBmatcher.cc:45-129 (Tracking::SearchLocalPoints) -----------------------

def test_oracle_kat_map_projection():
    """Radius 2.5 / 4.0 by viewing cosine times th times the level's scale; levels level-1 .. level; ratio test only when best
    and second-best lie on the same level; the stereo coordinate gate; occupied keypoints."""
    k2 = np.array([kp(100, 100, 1), kp(103, 100, 1), kp(103.5, 100, 0), kp(300, 300, 1)], dtype=O.KP_DTYPE)
    d2 = np.zeros((4, 32), np.uint8)
    d2[0, 0] = 0x03   # 2 bits
    d2[1, 0] = 0x07   # 3 bits
    d2[2, 0] = 0x0F   # 4 bits
    base = dict(level1=[1], valid1=[1], obs1=[1], mp_desc1=np.zeros((1, 32), np.uint8), k2=k2, d2=d2, u_right2=None,
                occupied2=np.zeros(4, np.uint8), bounds=BOUNDS, scale_factors=SF)
    proj = np.array([[101.0, 100.0, 90.0]], np.float32)
    # cos 0.999 -> r = 2.5 * 1.2 = 3.0: candidates 0 (dx 1), 1 (dx 2), 2 (dx 2.5).  best 2 bits, second 3 bits on the same level:
    # 2 > 0.6 * 3 -> rejected by the ratio test; with ratio 0.8 accepted
    assert O.search_by_projection_map(proj, [0.999], th=1.0, nnratio=0.6, **base)[0] == 0
    n, m21 = O.search_by_projection_map(proj, [0.999], th=1.0, nnratio=0.8, **base)
    assert n == 1 and list(m21) == [0, -1, -1, -1]
    # second-best on another level: no ratio test.  Occupy keypoint 1 -> second = keypoint 2 (octave 0)
    occ = np.array([0, 1, 0, 0], np.uint8)
    n, m21 = O.search_by_projection_map(proj, [0.999], th=1.0, nnratio=0.1, **{**base, "occupied2": occ})
    assert n == 1 and list(m21) == [0, -1, -1, -1]
    # a narrow window (level 0 point, r = 2.5) does not reach keypoint 1 / 2 and level 0 admits octave 0 (and -1) only
    n, m21 = O.search_by_projection_map(proj, [0.999], th=1.0, nnratio=0.8, **{**base, "level1": [0]})
    assert n == 0
    # stereo gate: |mTrackProjXR - mvuRight| > r rejects
    ur = np.array([95.0, -1, -1, -1], np.float32)
    n, m21 = O.search_by_projection_map(proj, [0.999], th=1.0, nnratio=0.9, **{**base, "u_right2": ur})
    assert list(m21) == [-1, 0, -1, -1]
    # th != 1 scales the radius; cos <= 0.998 -> 4.0
    far = np.array([[105.9, 100.0, 90.0]], np.float32)
    assert O.search_by_projection_map(far, [0.999], th=1.0, nnratio=0.9, **base)[0] == 1  # reaches keypoints 1, 2 only
    assert list(O.search_by_projection_map(far, [0.5], th=1.0, nnratio=0.9, **base)[1]) == [-1, 0, -1, -1]
    assert list(O.search_by_projection_map(far, [0.5], th=2.0, nnratio=0.9, **base)[1]) == [0, -1, -1, -1]


def map_scenario(seed, stereo):
    s = scenario(seed, stereo)
    rng = np.random.default_rng(1000 + seed)
    n1 = s["valid1"].size
    cam, z = s["cam"], s["x3Dc"][:, 2]
    u = (np.float32(cam[0]) * s["x3Dc"][:, 0] / z + np.float32(cam[2])).astype(np.float32)
    v = (np.float32(cam[1]) * s["x3Dc"][:, 1] / z + np.float32(cam[3])).astype(np.float32)
    proj = np.stack([u, v, (u - np.float32(s["mbf"]) / z).astype(np.float32)], axis=1).astype(np.float32)
    view_cos = rng.uniform(0.99, 1.0, n1).astype(np.float32)
    level = np.clip(s["octave1"] + rng.integers(0, 2, n1), 0, 7).astype(np.int32)
    valid = (s["valid1"] & (z > 0)).astype(np.uint8)
    return dict(proj=proj, view_cos=view_cos, level1=level, valid1=valid, obs1=s["obs1"], mp_desc1=s["mp_desc1"], k2=s["k2"],
                d2=s["d2"], u_right2=s["u_right2"], occupied2=s["occupied2"], bounds=s["bounds"], scale_factors=s["scale_factors"])


@pytest.mark.gpu
@pytest.mark.parametrize("seed,stereo,th,ratio", [(1, False, 1.0, 0.8), (2, True, 3.0, 0.8), (3, False, 5.0, 0.6), (4, True, 1.0, 0.9)])
def test_search_by_projection_map_gpu(seed, stereo, th, ratio):
    import orb_slam_cuda_b200 as X
    s = map_scenario(seed, stereo)
    n_ref, m_ref = O.search_by_projection_map(th=th, nnratio=ratio, **s)
    m = X.ORBmatcher(ratio, True)
    n, m21 = m.SearchByProjectionMap(s["proj"], s["view_cos"], s["level1"], s["valid1"], s["obs1"], s["mp_desc1"], s["k2"], s["d2"],
                                     s["u_right2"], s["occupied2"], s["bounds"], s["scale_factors"], th)
    assert n_ref > 30, "the scenario must produce matches"
    assert n == n_ref
    assert np.array_equal(m21, m_ref)


# ---- SearchByProjection(pKF, Scw, vpPoints, vpMatched, th) — src/ORBmatcher.cc:290-403 (loop detection) ------------------

def test_oracle_kat_sim3_projection():
    k2 = np.array([kp(100, 100, 1), kp(102, 100, 2), kp(101, 101, 0)], dtype=O.KP_DTYPE)
    d2 = np.zeros((3, 32), np.uint8)
    d2[0, 0], d2[1, 0], d2[2, 0] = 0x07, 0x01, 0x03
    uvr = np.array([[101.0, 100.0, 5.0], [101.0, 100.0, 5.0]], np.float32)
    mp = np.zeros((2, 32), np.uint8)
    # level 1 admits octaves 0..1: keypoint 2 (2 bits) beats keypoint 0 (3 bits); keypoint 1 (octave 2) is out although closest.
    # The second point finds keypoint 2 taken (vpMatched set) and falls back to keypoint 0.
    n, m21 = O.search_by_projection_sim3(uvr, [1, 1], [1, 1], mp, k2, d2, [0, 0, 0], BOUNDS)
    assert n == 2 and list(m21) == [1, -1, 0]
    n, m21 = O.search_by_projection_sim3(uvr, [2, 2], [1, 0], mp, k2, d2, [0, 0, 0], BOUNDS)
    assert n == 1 and list(m21) == [-1, 0, -1]
    far = mp.copy()
    far[:, :7] = 0xFF  # 56 - 3 bits > TH_LOW
    assert O.search_by_projection_sim3(uvr, [1, 1], [1, 1], far, k2, d2, [0, 0, 0], BOUNDS)[0] == 0


@pytest.mark.gpu
@pytest.mark.parametrize("seed,th", [(11, 10), (12, 4)])
def test_search_by_projection_sim3_gpu(seed, th):
    import orb_slam_cuda_b200 as X
    s = map_scenario(seed, False)
    rng = np.random.default_rng(seed)
    uvr = s["proj"].copy()
    uvr[:, 2] = (np.float32(th) * s["scale_factors"][s["level1"]]).astype(np.float32)
    occ = (rng.random(s["k2"].size) < 0.15).astype(np.uint8)
    n_ref, m_ref = O.search_by_projection_sim3(uvr, s["level1"], s["valid1"], s["mp_desc1"], s["k2"], s["d2"], occ, s["bounds"], 50)
    n, m21 = X.ORBmatcher(0.75, True).SearchByProjectionSim3(uvr, s["level1"], s["valid1"], s["mp_desc1"], s["k2"], s["d2"], occ,
                                                             s["bounds"], 50)
    assert n_ref > 50 and n == n_ref and np.array_equal(m21, m_ref)


@pytest.mark.gpu
@pytest.mark.parametrize("seed,th", [(21, 30.0), (22, 60.0)])
def test_search_by_projection_dense_competition(seed, th):
    """Few distinct descriptors and wide windows: many points want the same keypoint, best candidates are taken by earlier
    points — conflict rounds, exhausted candidate caches and rescans of the replay kernels (frame-to-frame, local map, Sim3)."""
    import orb_slam_cuda_b200 as X
    s = scenario(seed, True)
    rng = np.random.default_rng(seed)
    base = rng.integers(0, 256, (10, 32), dtype=np.uint8)
    n1, n2 = s["valid1"].size, s["k2"].size
    mp = base[rng.integers(0, 10, n1)].copy()
    d2 = base[rng.integers(0, 10, n2)].copy()
    mp[:, 3] ^= rng.integers(0, 8, n1).astype(np.uint8)
    d2[:, 7] ^= rng.integers(0, 8, n2).astype(np.uint8)
    s["mp_desc1"], s["d2"] = mp, d2
    n_ref, m_ref = O.search_by_projection_frame(s["x3Dc"], s["valid1"], s["obs1"], mp, s["octave1"], s["angle1"], s["k2"], d2,
                                                s["u_right2"], s["occupied2"], s["cam"], s["mbf"], s["bounds"], s["scale_factors"], th, 0,
                                                True)
    m = X.ORBmatcher(0.9, True)
    n, m21 = m.SearchByProjection(s["x3Dc"], s["valid1"], s["obs1"], mp, s["octave1"], s["angle1"], s["k2"], d2, s["u_right2"],
                                  s["occupied2"], s["cam"], s["mbf"], s["bounds"], s["scale_factors"], th, 0)
    assert n_ref > 100 and n == n_ref and np.array_equal(m21, m_ref)
    # local map and Sim3 variants on the same crowded data
    ms = map_scenario(seed, True)
    ms["mp_desc1"] = base[rng.integers(0, 10, ms["valid1"].size)].copy()
    ms["d2"] = base[rng.integers(0, 10, ms["k2"].size)].copy()
    ms["d2"][:, 5] ^= rng.integers(0, 4, ms["k2"].size).astype(np.uint8)
    occ = np.zeros(ms["k2"].size, np.uint8)
    r = O.search_by_projection_map(ms["proj"], ms["view_cos"], ms["level1"], ms["valid1"], ms["obs1"], ms["mp_desc1"], ms["k2"], ms["d2"],
                                   ms["u_right2"], occ, ms["bounds"], ms["scale_factors"], th / 3, 0.95)
    g = X.ORBmatcher(0.95, True).SearchByProjectionMap(ms["proj"], ms["view_cos"], ms["level1"], ms["valid1"], ms["obs1"], ms["mp_desc1"],
                                                       ms["k2"], ms["d2"], ms["u_right2"], occ, ms["bounds"], ms["scale_factors"], th / 3)
    assert r[0] > 50 and g[0] == r[0] and np.array_equal(g[1], r[1])
    uvr = ms["proj"].copy()
    uvr[:, 2] = (np.float32(th) * ms["scale_factors"][ms["level1"]]).astype(np.float32)
    r = O.search_by_projection_sim3(uvr, ms["level1"], ms["valid1"], ms["mp_desc1"], ms["k2"], ms["d2"], occ, ms["bounds"], 50)
    g = m.SearchByProjectionSim3(uvr, ms["level1"], ms["valid1"], ms["mp_desc1"], ms["k2"], ms["d2"], occ, ms["bounds"], 50)
    assert r[0] > 50 and g[0] == r[0] and np.array_equal(g[1], r[1])


def test_oracle_kat_fuse_match():
    """Fuse's candidate loop (src/ORBmatcher.cc:886-945): level gate, chi-square gate (mono 5.99 / stereo 7.8 on e2 * invSigma2), best
    distance with the first candidate winning ties; the Sim3 overload (:1049-1075) has no gate."""
    k2 = np.array([kp(200, 200, 1), kp(202, 200, 1), kp(201, 201, 3), kp(210, 200, 1)], dtype=O.KP_DTYPE)
    d2 = np.zeros((4, 32), np.uint8)
    d2[0, 0], d2[1, 0], d2[2, 0], d2[3, 0] = 0x03, 0x0C, 0x00, 0x00       # 2, 2, 0, 0 bits from the zero descriptor
    inv = (1.0 / (SF * SF)).astype(np.float32)
    uvr = np.array([[200.5, 200.0, 20.0]], np.float32)
    mp = np.zeros((1, 32), np.uint8)
    # level 3 is outside [pred - 1, pred] = [0, 1]; keypoint 3 is 9.5 px away: e2 * inv = 90 / 1.44 > 5.99 -> gated; 0 beats 1 (first)
    idx, dist = O.fuse_match(uvr, [190.0], [1], [1], mp, k2, d2, None, inv, BOUNDS)
    assert (idx[0], dist[0]) == (0, 2)
    # without the gate (Sim3 overload) the far keypoint with distance 0 wins
    idx, dist = O.fuse_match(uvr, None, [1], [1], mp, k2, d2, None, inv, BOUNDS)
    assert (idx[0], dist[0]) == (3, 0)
    # stereo keypoints add the right-image error: er = 190 - 150 = 40 -> gated, so keypoint 1 (mono) is left
    ur2 = np.array([150.0, -1.0, -1.0, -1.0], np.float32)
    idx, dist = O.fuse_match(uvr, [190.0], [1], [1], mp, k2, d2, ur2, inv, BOUNDS)
    assert (idx[0], dist[0]) == (1, 2)
    assert O.fuse_match(uvr, [190.0], [1], [0], mp, k2, d2, None, inv, BOUNDS)[0][0] == -1
    assert tuple(a[0] for a in O.fuse_match(np.array([[50.0, 50.0, 5.0]], np.float32), None, [1], [1], mp, k2, d2, None, inv, BOUNDS)) == (-1, 256)


@pytest.mark.gpu
@pytest.mark.parametrize("seed,stereo,gate,th", [(31, True, True, 3.0), (32, False, True, 4.0), (33, True, False, 4.0), (34, False, False, 10.0)])
def test_fuse_match_gpu(seed, stereo, gate, th):
    import orb_slam_cuda_b200 as X
    s = map_scenario(seed, stereo)
    uvr = s["proj"].copy()
    uvr[:, 2] = (np.float32(th) * s["scale_factors"][s["level1"]]).astype(np.float32)
    ur1 = s["proj"][:, 2].copy() if gate else None
    inv = (np.float32(1.0) / (s["scale_factors"] * s["scale_factors"])).astype(np.float32)
    r_idx, r_dist = O.fuse_match(uvr, ur1, s["level1"], s["valid1"], s["mp_desc1"], s["k2"], s["d2"], s["u_right2"], inv, s["bounds"])
    idx, dist = X.ORBmatcher().FuseMatch(uvr, ur1, s["level1"], s["valid1"], s["mp_desc1"], s["k2"], s["d2"], s["u_right2"], inv, s["bounds"])
    assert (r_idx >= 0).sum() > 50 and np.array_equal(idx, r_idx) and np.array_equal(dist, r_dist)
    assert ((r_dist <= 50) & (r_idx >= 0)).sum() > 20


# ---- orbx_frame: the current frame resident on the device (Frame::AssignFeaturesToGrid, src/Frame.cc:229-244) -------------------------

@pytest.mark.gpu
@pytest.mark.parametrize("seed,stereo", [(1, False), (3, True)])
def test_device_frame_handle(seed, stereo):
    """The grid binned on the device equals the reference's mGrid (cell contents in push_back order = ascending keypoint index:
    checked through the oracle's GetFeaturesInArea, which tests/test_oracle_ref_matcher.py pins to the reference's Frame.cc), and the
    handle entries of the two per-frame tracking matchers return exactly what the host-array entries return."""
    import orb_slam_cuda_b200 as X
    s = scenario(seed, stereo)
    k2, n2 = s["k2"], s["k2"].size
    fr = X.DeviceFrame(k2, s["d2"], s["bounds"], s["u_right2"])
    cs, so = fr.grid()
    assert cs[0] == 0 and cs[-1] == n2 and sorted(so.tolist()) == list(range(n2))
    w, h = s["bounds"][1], s["bounds"][3]
    for cell in (0, 64 * 48 // 2 + 7, 1000, 64 * 48 - 1):  # a cell = everything GetFeaturesInArea finds with the cell's box as the window
        inside = so[cs[cell]:cs[cell + 1]]
        assert list(inside) == sorted(inside)
        ix, iy = cell // 48, cell % 48
        px = np.floor((k2["x"] * (np.float32(64) / np.float32(w))).astype(np.float32) + np.float32(0.5)).astype(int)  # PosInGrid, src/Frame.cc:381-391
        py = np.floor((k2["y"] * (np.float32(48) / np.float32(h))).astype(np.float32) + np.float32(0.5)).astype(int)
        assert list(inside) == list(np.flatnonzero((px == ix) & (py == iy)))
    m = X.ORBmatcher(0.9, True)
    a = (s["x3Dc"], s["valid1"], s["obs1"], s["mp_desc1"], s["octave1"], s["angle1"])
    n_h, m_h = m.SearchByProjection(*a, k2, s["d2"], s["u_right2"], s["occupied2"], s["cam"], s["mbf"], s["bounds"], s["scale_factors"], 15.0, 0)
    n_f, m_f = m.SearchByProjectionFrame(*a, fr, s["occupied2"], s["cam"], s["mbf"], s["scale_factors"], 15.0, 0)
    n_o, m_o = O.search_by_projection_frame(*a, k2, s["d2"], s["u_right2"], s["occupied2"], s["cam"], s["mbf"], s["bounds"], s["scale_factors"], 15.0, 0, True)
    assert n_f == n_h == n_o > 50 and np.array_equal(m_f, m_h) and np.array_equal(m_f, m_o)
    ms = map_scenario(seed, stereo)
    mm = X.ORBmatcher(0.8, True)
    b = (ms["proj"], ms["view_cos"], ms["level1"], ms["valid1"], ms["obs1"], ms["mp_desc1"])
    n_h, m_h = mm.SearchByProjectionMap(*b, k2, ms["d2"], ms["u_right2"], ms["occupied2"], ms["bounds"], ms["scale_factors"], 3.0)
    n_f, m_f = mm.SearchByProjectionMapFrame(*b, fr, ms["occupied2"], ms["scale_factors"], 3.0)
    assert n_f == n_h > 50 and np.array_equal(m_f, m_h)
    # an empty frame works too
    e = X.DeviceFrame(k2[:0], s["d2"][:0], s["bounds"])
    n_e, m_e = m.SearchByProjectionFrame(*a, e, np.zeros(0, np.uint8), s["cam"], s["mbf"], s["scale_factors"], 15.0, 0)
    assert n_e == 0
