"""Parity of the CUDA matchers (through the C ABI) with the CPU oracle: DescriptorDistance, brute force,
SearchForInitialization, ComputeStereoMatches, all-vs-all keyframe scoring.

Bars (BASELINE.json north_star): match indices and distances bit-exact; the float outputs of the stereo matcher
(mvuRight, mvDepth) are compared bit for bit as well (tolerance 0)."""
import os

import numpy as np
import pytest

from oracle import oracle_lib as O
from orb_slam_cuda_b200 import synth
import orb_slam_cuda_b200 as X

pytestmark = pytest.mark.gpu


def test_descriptor_distance_kat():
    z, o = np.zeros(32, np.uint8), np.full(32, 255, np.uint8)
    assert X.ORBmatcher.DescriptorDistance(z, o) == 256  # SURVEY.md 8c known answer
    assert X.ORBmatcher.DescriptorDistance(z, z) == 0
    rng = np.random.default_rng(1)
    a = rng.integers(0, 256, (5000, 32), dtype=np.uint8)
    b = rng.integers(0, 256, (5000, 32), dtype=np.uint8)
    got = X.ORBmatcher.DescriptorDistance(a, b)
    want = np.unpackbits(a ^ b, axis=1).sum(1)
    assert np.array_equal(got, want)
    for i in range(0, 5000, 977):
        assert O.descriptor_distance(a[i], b[i]) == got[i]


@pytest.mark.parametrize("nq,nd", [(2000, 2000), (1, 1), (37, 5), (5, 3001), (1300, 257), (129, 0)])
def test_bruteforce(nq, nd):
    q, d = synth.descriptor_sets(100 + nq + nd, nq, max(nd, 1))
    d = d[:nd]
    m = X.ORBmatcher(0.6)
    got = m.match_bruteforce(q, d, th=50)
    want = O.match_bruteforce(q, d, 50, 0.6)
    for g, w, name in zip(got, want, ["idx", "d1", "d2", "match"]):
        assert np.array_equal(g, w), f"{name}: {(g != w).sum()} of {nq} differ"
    if nd >= 100:
        assert (got[3] >= 0).sum() > 0 and (got[3] < 0).sum() > 0  # both branches of the test fire


def test_bruteforce_ties():
    # duplicates in the database: the first index wins and second == best kills the ratio test
    rng = np.random.default_rng(5)
    d = rng.integers(0, 256, (64, 32), dtype=np.uint8)
    d[40] = d[7]
    q = d[[7, 3]].copy()
    got = X.ORBmatcher(0.6).match_bruteforce(q, d)
    want = O.match_bruteforce(q, d, 50, 0.6)
    for g, w in zip(got, want):
        assert np.array_equal(g, w)
    assert got[0][0] == 7 and got[1][0] == 0 and got[2][0] == 0 and got[3][0] == -1 and got[3][1] == 3


def _extract_pair(kind, k, nfeat, second):
    ox = O.Extractor(nfeat, 1.2, 8, 20, 7)
    a = synth.frame(kind, k)
    k1, d1 = ox.extract(a)
    k2, d2 = ox.extract(second(a))
    return a.shape, k1, d1, k2, d2


@pytest.mark.parametrize("kind,nfeat,window,ratio,ori", [("kitti", 2000, 100, 0.9, True), ("tum", 1000, 100, 0.9, True),
                                                          ("tum", 1000, 10, 0.9, False), ("euroc", 1200, 30, 0.6, True)])
def test_search_for_initialization(kind, nfeat, window, ratio, ori):
    shape, k1, d1, k2, d2 = _extract_pair(kind, 31, nfeat, lambda a: synth.shifted_frame(kind, 31))
    h, w = shape
    bounds = (0.0, float(w), 0.0, float(h))
    prev = np.stack([k1["x"], k1["y"]], 1).astype(np.float32)
    n_ref, m_ref, prev_ref = O.search_for_initialization(k1, d1, k2, d2, bounds, prev, window, ratio, ori)
    prev_gpu = prev.copy()
    n_gpu, m_gpu = X.ORBmatcher(ratio, ori).SearchForInitialization(k1, d1, k2, d2, bounds, prev_gpu, window)
    assert n_gpu == n_ref
    assert np.array_equal(m_gpu, m_ref), f"{(m_gpu != m_ref).sum()} of {len(m_ref)} matches differ"
    assert np.array_equal(prev_gpu.view(np.uint32), prev_ref.reshape(-1, 2).view(np.uint32))
    if window >= 30:
        assert n_ref > 20  # the case exercises real matches
    # second round with the updated vbPrevMatched (Tracking calls it again on the next frame, src/Tracking.cc:671)
    n2_ref, m2_ref, _ = O.search_for_initialization(k1, d1, k2, d2, bounds, prev_ref, window, ratio, ori)
    n2_gpu, m2_gpu = X.ORBmatcher(ratio, ori).SearchForInitialization(k1, d1, k2, d2, bounds, prev_gpu, window)
    assert n2_gpu == n2_ref and np.array_equal(m2_gpu, m2_ref)


def test_search_for_initialization_random_descriptors():
    # dense conflicts: few distinct descriptors so stealing and the vMatchedDistance skip fire often
    rng = np.random.default_rng(9)
    n = 600
    k = np.zeros(n, X.KP_DTYPE)
    k["x"], k["y"] = rng.uniform(0, 640, n), rng.uniform(0, 480, n)
    k["angle"] = rng.uniform(0, 360, n)
    k["octave"] = rng.integers(0, 3, n)
    k2 = k.copy()
    k2["x"] += rng.normal(0, 3, n).astype(np.float32)
    k2["angle"] = (k["angle"] + rng.normal(0, 20, n)).astype(np.float32) % 360
    base = rng.integers(0, 256, (20, 32), dtype=np.uint8)
    d1 = base[rng.integers(0, 20, n)].copy()
    d2 = base[rng.integers(0, 20, n)].copy()
    flips = rng.integers(0, 256, (n, 32), dtype=np.uint8) & rng.integers(0, 256, (n, 32), dtype=np.uint8) & \
        rng.integers(0, 256, (n, 32), dtype=np.uint8) & rng.integers(0, 256, (n, 32), dtype=np.uint8)
    d2 ^= flips
    bounds = (0.0, 640.0, 0.0, 480.0)
    prev = np.stack([k["x"], k["y"]], 1).astype(np.float32)
    for ratio in (0.9, 0.99):
        n_ref, m_ref, prev_ref = O.search_for_initialization(k, d1, k2, d2, bounds, prev, 60, ratio, True)
        pg = prev.copy()
        n_gpu, m_gpu = X.ORBmatcher(ratio, True).SearchForInitialization(k, d1, k2, d2, bounds, pg, 60)
        assert n_gpu == n_ref and np.array_equal(m_gpu, m_ref)
        assert np.array_equal(pg.view(np.uint32), prev_ref.reshape(-1, 2).view(np.uint32))


def test_stereo_matches():
    # EuRoC stereo (Examples/Stereo/EuRoC.yaml: 752x480, 1200 features per eye, bf 47.906, fx 435.2)
    left, right = synth.stereo_pair("euroc", 5)
    bf, fx = 47.90639384423901, 435.2046959714599
    oxL, oxR = O.Extractor(1200, 1.2, 8, 20, 7), O.Extractor(1200, 1.2, 8, 20, 7)
    kL, dL = oxL.extract(left)
    kR, dR = oxR.extract(right)
    u_ref, z_ref = O.stereo_matches(oxL, oxR, kL, dL, kR, dR, bf, fx)
    gxL, gxR = X.ORBextractor(1200, 1.2, 8, 20, 7), X.ORBextractor(1200, 1.2, 8, 20, 7)
    gkL, gdL = gxL(left)
    gkR, gdR = gxR(right)
    assert np.array_equal(gkL.view(np.uint8), kL.view(np.uint8)) and np.array_equal(gdR, dR)
    u, z = X.ComputeStereoMatches(gxL, gxR, gkL, gdL, gkR, gdR, bf, fx)
    assert (u_ref >= 0).sum() > 100, "the synthetic pair should yield stereo matches"
    assert np.array_equal(u.view(np.uint32), u_ref.view(np.uint32)), f"{(u != u_ref).sum()} of {len(u)} uRight differ"
    assert np.array_equal(z.view(np.uint32), z_ref.view(np.uint32))


def test_stereo_matches_kitti():
    left, right = synth.stereo_pair("kitti", 2, max_disp=60)
    bf, fx = 386.1448, 718.856
    oxL, oxR = O.Extractor(2000, 1.2, 8, 20, 7), O.Extractor(2000, 1.2, 8, 20, 7)
    kL, dL = oxL.extract(left)
    kR, dR = oxR.extract(right)
    u_ref, z_ref = O.stereo_matches(oxL, oxR, kL, dL, kR, dR, bf, fx)
    gxL, gxR = X.ORBextractor(2000, 1.2, 8, 20, 7), X.ORBextractor(2000, 1.2, 8, 20, 7)
    gkL, gdL = gxL(left)
    gkR, gdR = gxR(right)
    u, z = X.ComputeStereoMatches(gxL, gxR, gkL, gdL, gkR, gdR, bf, fx)
    assert np.array_equal(u.view(np.uint32), u_ref.view(np.uint32)) and np.array_equal(z.view(np.uint32), z_ref.view(np.uint32))


def test_stereo_matches_on_device_resident_output():
    """orbx_stereo_match_last: ComputeStereoMatches on what the two extractors left on the device = the host-array entry point,
    bit for bit; refused after any call that is not a single-frame extract"""
    for kind, nf, bf, fx, kw in (("euroc", 1200, 47.90639384423901, 435.2046959714599, {}), ("kitti", 2000, 386.1448, 718.856, dict(max_disp=60))):
        left, right = synth.stereo_pair(kind, 7, **kw)
        gxL, gxR = X.ORBextractor(nf, 1.2, 8, 20, 7), X.ORBextractor(nf, 1.2, 8, 20, 7)
        for rep in range(3):  # the second and third call replay the captured graph
            gkL, gdL = gxL(left)
            gkR, gdR = gxR(right)
            u_ref, z_ref = X.ComputeStereoMatches(gxL, gxR, gkL, gdL, gkR, gdR, bf, fx)
            u, z = X.ComputeStereoMatchesLast(gxL, gxR, len(gkL), bf, fx)
            assert (u_ref >= 0).sum() > 100
            assert np.array_equal(u.view(np.uint32), u_ref.view(np.uint32)) and np.array_equal(z.view(np.uint32), z_ref.view(np.uint32))
        with pytest.raises(X.OrbxError):  # output capacity below the left keypoint count
            X.ComputeStereoMatchesLast(gxL, gxR, len(gkL) - 1, bf, fx)
        gxL.extract_batch(np.stack([left, left]))  # a batch call: the single-frame output is gone
        with pytest.raises(X.OrbxError):
            X.ComputeStereoMatchesLast(gxL, gxR, len(gkL), bf, fx)
        gxL(left)
        u, z = X.ComputeStereoMatchesLast(gxL, gxR, len(gkL), bf, fx)
        assert np.array_equal(u.view(np.uint32), u_ref.view(np.uint32))


def test_allpairs_scores():
    import torch
    rng = np.random.default_rng(3)
    kq, kd, n = 3, 4, 700
    db = rng.integers(0, 256, (kd, n, 32), dtype=np.uint8)
    q = rng.integers(0, 256, (kq, n, 32), dtype=np.uint8)
    # query keyframe i shares noisy copies of a third of database keyframe (i + 1) % kd
    for i in range(kq):
        src = db[(i + 1) % kd, : n // 3].copy()
        noise = (rng.integers(0, 256, src.shape, dtype=np.uint8) & rng.integers(0, 256, src.shape, dtype=np.uint8)
                 & rng.integers(0, 256, src.shape, dtype=np.uint8) & rng.integers(0, 256, src.shape, dtype=np.uint8))
        q[i, : n // 3] = src ^ noise
    want = np.zeros((kq, kd), np.int32)
    for i in range(kq):
        for j in range(kd):
            want[i, j] = (O.match_bruteforce(q[i], db[j], 50, 0.6)[3] >= 0).sum()
    dq, ddb = torch.from_numpy(q).cuda(), torch.from_numpy(db).cuda()
    score = torch.full((kq, kd), -1, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()  # the matcher launches on its own non-blocking stream
    X.ORBmatcher(0.6).match_allpairs_device(dq.data_ptr(), kq, ddb.data_ptr(), kd, n, score.data_ptr(),
                                            stream=torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert np.array_equal(score.cpu().numpy(), want)
    assert want.max() > 50


def test_allpairs_sharded_entry_single_rank():
    """orbx_allpairs_sharded with world = 1 (no communicator) is the plain all-vs-all kernel; the column-mapped form of the
    kernel (strided score rows) is what the multi-GPU path launches — tools/multi_gpu_check.py covers world > 1."""
    import torch
    from orb_slam_cuda_b200 import sharding
    rng = np.random.default_rng(5)
    k, n = 9, 700
    base = rng.integers(0, 256, size=(3, n, 32), dtype=np.uint8)
    desc = np.stack([np.roll(base[i % 3], i, axis=0) ^ (rng.random((n, 32)) < 0.02).astype(np.uint8) for i in range(k)])
    d = torch.from_numpy(desc).cuda()
    want = torch.empty((k, k), dtype=torch.int32, device="cuda")
    m = X.ORBmatcher(0.6, True)
    m.match_allpairs_device(d.data_ptr(), k, d.data_ptr(), k, n, want.data_ptr())
    comm = sharding.NcclComm(torch.device("cuda", 0))
    got, moved = sharding.allpairs_sharded_native(comm, d, k)
    torch.cuda.synchronize()
    assert moved == 0 and torch.equal(got, want) and int((want > 0).sum()) >= k
    for qi, dj in [(0, 0), (4, 7), (8, 2)]:
        _, _, _, mt = O.match_bruteforce(desc[qi], desc[dj], 50, 0.6)
        assert int((mt >= 0).sum()) == int(want[qi, dj])


def test_allpairs_sharded_multi_gpu():
    """world = 2 under torchrun when the box has two GPUs (the 1-GPU box skips; tools/gpu_run.sh multi runs it on N GPUs)."""
    import subprocess, sys, torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    root = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                        "--master-port", "29533", os.path.join(root, "tools", "multi_gpu_check.py")], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
