"""ORBmatcher::SearchByBoW, both overloads — reference src/ORBmatcher.cc:159-288 (KeyFrame vs Frame) and :522-655 (KeyFrame vs
KeyFrame); SURVEY.md §8f rank 2.  CPU: known-answer cases for the oracle's reading of the loop (a taken feature is skipped by
later iterations, best/second over the free ones, second defaults to 256, `<=` vs `<` TH_LOW, rotation histogram).  GPU: the CUDA
path through the C ABI against the oracle, bit-exact, on descriptors of real extractions grouped by a toy vocabulary."""
import numpy as np
import pytest

from oracle import oracle_lib as O


def feature_vectors(desc, bits=8):
    """A stand-in for DBoW2's FeatureVector: node id = the first `bits` descriptor bits of byte 5.., features in index order
    (DBoW2 appends feature i to its node's vector while walking i = 0 .. N-1)."""
    node = (desc[:, 5].astype(np.int32) | (desc[:, 17].astype(np.int32) << 8)) & ((1 << bits) - 1)
    return {int(k): np.flatnonzero(node == k).astype(np.uint32) for k in np.unique(node)}


def common_nodes(fv1, fv2):
    """The merge of :180-256: nodes present in both maps, ascending."""
    nodes = sorted(set(fv1) & set(fv2))
    s1, s2 = [0], [0]
    for k in nodes:
        s1.append(s1[-1] + fv1[k].size)
        s2.append(s2[-1] + fv2[k].size)
    cat = lambda fv: np.concatenate([fv[k] for k in nodes]) if nodes else np.zeros(0, np.uint32)
    return np.array(s1, np.int32), cat(fv1), np.array(s2, np.int32), cat(fv2)


def test_oracle_kat_taken_ratio_and_threshold():
    d1 = np.zeros((3, 32), np.uint8)
    d2 = np.zeros((3, 32), np.uint8)
    d2[0, 0] = 0x01                    # 1 bit from d1[*]
    d2[1, 0:2] = 0xFF                  # 16 bits
    d2[2, 0:5] = 0xFF                  # 40 bits
    one3 = np.ones(3, np.uint8)
    ang = np.zeros(3, np.float32)
    s1, i1, s2, i2 = [0, 3], [0, 1, 2], [0, 3], [0, 1, 2]
    # feature 0 takes d2[0] (1 < 0.7 * 16); feature 1 then sees {16, 40}: 16 < 28 -> takes d2[1]; feature 2 sees only {40}:
    # second defaults to 256, 40 <= 50 and 40 < 179.2 -> takes d2[2]
    n, m12, m21 = O.search_by_bow(s1, i1, s2, i2, one3, d1, ang, one3, d2, ang, 50, 0.7, False)
    assert n == 3 and list(m12) == [0, 1, 2] and list(m21) == [0, 1, 2]
    # KeyFrame overload is strict: with th_dist = 39 the last one fails, with 40 it passes
    assert O.search_by_bow(s1, i1, s2, i2, one3, d1, ang, one3, d2, ang, 39, 0.7, False)[0] == 2
    assert O.search_by_bow(s1, i1, s2, i2, one3, d1, ang, one3, d2, ang, 40, 0.7, False)[0] == 3
    # a tighter ratio rejects feature 1 (16 !< 0.3 * 40) and then feature 2 faces {16, 40}: 16 !< 12 -> nothing
    n, m12, m21 = O.search_by_bow(s1, i1, s2, i2, one3, d1, ang, one3, d2, ang, 50, 0.3, False)
    assert n == 1 and list(m12) == [0, -1, -1] and list(m21) == [0, -1, -1]
    # invalid first-argument features are skipped; invalid second-argument features are never candidates
    n, m12, m21 = O.search_by_bow(s1, i1, s2, i2, [0, 1, 1], d1, ang, [1, 0, 1], d2, ang, 50, 0.7, False)
    assert n == 2 and list(m12) == [-1, 0, 2] and list(m21) == [1, -1, 2]
    # features in different nodes never meet
    n, m12, m21 = O.search_by_bow([0, 1, 3], [0, 1, 2], [0, 2, 3], [1, 2, 0], one3, d1, ang, one3, d2, ang, 50, 0.7, False)
    assert n == 2 and list(m12) == [1, 0, -1]


def test_oracle_kat_rotation_histogram():
    n = 12
    d1 = np.zeros((n, 32), np.uint8)
    d1[np.arange(n), 8 + np.arange(n)] = 0xFF  # distinct descriptors, 16 bits apart from one another
    d2 = d1.copy()
    a1 = np.full(n, 90.0, np.float32)
    a2 = np.full(n, 80.0, np.float32)
    a2[3] = 300.0   # rot = -210 + 360 = 150 -> bin 5; the rest rot = 10 -> bin 0
    one = np.ones(n, np.uint8)
    s, i = [0, n], np.arange(n)
    cnt, m12, m21 = O.search_by_bow(s, i, s, i, one, d1, a1, one, d2, a2, 50, 0.7, True)
    assert cnt == n - 1 and m12[3] == -2 and m21[3] == -2 and all(m12[k] == k for k in range(n) if k != 3)
    cnt, m12, _ = O.search_by_bow(s, i, s, i, one, d1, a1, one, d2, a2, 50, 0.7, False)
    assert cnt == n and m12[3] == 3


def bow_scenario(seed, kf_kf):
    from orb_slam_cuda_b200 import synth
    rng = np.random.default_rng(seed)
    img = synth.frame("tum", seed)
    cur = np.roll(np.roll(img, int(rng.integers(-3, 4)), axis=0), int(rng.integers(2, 6)), axis=1)
    ox = O.Extractor(1000, 1.2, 8, 20, 7)
    k1, d1 = ox.extract(img)
    k1, d1 = k1.copy(), d1.copy()
    k2, d2 = ox.extract(cur)
    k2, d2 = k2.copy(), d2.copy()
    s1, i1, s2, i2 = common_nodes(feature_vectors(d1, 5), feature_vectors(d2, 5))
    valid1 = (rng.random(k1.size) < 0.8).astype(np.uint8)
    valid2 = (rng.random(k2.size) < 0.8).astype(np.uint8) if kf_kf else np.ones(k2.size, np.uint8)
    a2 = k2["angle"].copy()
    twist = rng.random(k2.size) < 0.1
    a2[twist] = (a2[twist] + rng.uniform(40, 300, int(twist.sum()))).astype(np.float32) % np.float32(360)
    return dict(start1=s1, idx1=i1, start2=s2, idx2=i2, valid1=valid1, desc1=d1, angle1=k1["angle"], valid2=valid2, desc2=d2, angle2=a2)


def test_oracle_matches_python_loop():
    """The oracle (C++) against a line-by-line Python restatement of :176-285 on a real scenario."""
    s = bow_scenario(3, True)
    d1, d2 = s["desc1"], s["desc2"]
    n1, n2 = s["valid1"].size, s["valid2"].size
    m12, taken, hist = [-1] * n1, [False] * n2, [[] for _ in range(30)]
    cnt = 0
    for k in range(s["start1"].size - 1):
        for a in s["idx1"][s["start1"][k]:s["start1"][k + 1]]:
            if not s["valid1"][a]:
                continue
            b1, b2, bi = 256, 256, -1
            for b in s["idx2"][s["start2"][k]:s["start2"][k + 1]]:
                if taken[b] or not s["valid2"][b]:
                    continue
                dist = int(np.unpackbits(d1[a] ^ d2[b]).sum())
                if dist < b1:
                    b2, b1, bi = b1, dist, int(b)
                elif dist < b2:
                    b2 = dist
            if b1 < 50 and np.float32(b1) < np.float32(0.75) * np.float32(b2):
                m12[a], taken[bi] = bi, True
                rot = np.float32(s["angle1"][a]) - np.float32(s["angle2"][bi])
                if rot < 0:
                    rot = np.float32(rot + np.float32(360))
                v = float(np.float32(rot * np.float32(1.0 / 30.0)))
                bn = int(np.floor(v + 0.5)) if v >= 0 else int(np.ceil(v - 0.5))
                hist[0 if bn == 30 else bn].append(int(a))
                cnt += 1
    order = sorted(range(30), key=lambda b: (-len(hist[b]), b))
    keep = {order[0]}
    if len(hist[order[1]]) >= 0.1 * len(hist[order[0]]):
        keep.add(order[1])
        if len(hist[order[2]]) >= 0.1 * len(hist[order[0]]):
            keep.add(order[2])
    for b in range(30):
        if b not in keep:
            for a in hist[b]:
                m12[a] = -2
                cnt -= 1
    n, o12, _ = O.search_by_bow(th_dist=49, nnratio=0.75, check_ori=True, **{("d1" if k == "desc1" else "d2" if k == "desc2" else k): v
                                                                             for k, v in s.items()})
    assert cnt > 100 and n == cnt and list(o12) == m12


@pytest.mark.gpu
@pytest.mark.parametrize("seed,kf_kf,ratio,ori", [(1, False, 0.7, True), (2, False, 0.75, True), (3, True, 0.75, True),
                                                  (4, True, 0.75, False), (5, False, 0.9, True)])
def test_search_by_bow_gpu(seed, kf_kf, ratio, ori):
    import orb_slam_cuda_b200 as X
    s = bow_scenario(seed, kf_kf)
    th = 49 if kf_kf else 50
    n_ref, r12, r21 = O.search_by_bow(s["start1"], s["idx1"], s["start2"], s["idx2"], s["valid1"], s["desc1"], s["angle1"], s["valid2"],
                                      s["desc2"], s["angle2"], th, ratio, ori)
    n, m12, m21 = X.ORBmatcher(ratio, ori).SearchByBoW(th_dist=th, **s)
    assert n_ref > 100 and n == n_ref and np.array_equal(m12, r12) and np.array_equal(m21, r21)
    if ori:
        assert (r12 == -2).sum() > 0  # the histogram removed something, so that path is exercised


@pytest.mark.gpu
def test_search_by_bow_one_big_node_and_empty():
    """All features in one node (a vocabulary level near the root): every item scans every feature of the other image."""
    import orb_slam_cuda_b200 as X
    s = bow_scenario(7, False)
    n1, n2 = s["valid1"].size, s["valid2"].size
    s.update(start1=np.array([0, n1], np.int32), idx1=np.arange(n1, dtype=np.uint32), start2=np.array([0, n2], np.int32),
             idx2=np.arange(n2, dtype=np.uint32))
    ref = O.search_by_bow(s["start1"], s["idx1"], s["start2"], s["idx2"], s["valid1"], s["desc1"], s["angle1"], s["valid2"], s["desc2"],
                          s["angle2"], 50, 0.7, True)
    m = X.ORBmatcher(0.7, True)
    got = m.SearchByBoW(**s)
    assert ref[0] > 100 and got[0] == ref[0] and np.array_equal(got[1], ref[1]) and np.array_equal(got[2], ref[2])
    # no common node / nothing valid
    e = np.zeros(0, np.uint32)
    n, m12, m21 = m.SearchByBoW(np.zeros(1, np.int32), e, np.zeros(1, np.int32), e, s["valid1"], s["desc1"], s["angle1"], s["valid2"],
                                s["desc2"], s["angle2"])
    assert n == 0 and (m12 == -1).all() and (m21 == -1).all()
    s["valid1"] = np.zeros(n1, np.uint8)
    n, m12, m21 = m.SearchByBoW(**s)
    assert n == 0 and (m12 == -1).all() and (m21 == -1).all()
    with pytest.raises(X.OrbxError):
        m.SearchByBoW(np.array([0, 1], np.int32), np.array([n1], np.uint32), np.array([0, 1], np.int32), np.array([0], np.uint32),
                      np.ones(n1, np.uint8), s["desc1"], s["angle1"], s["valid2"], s["desc2"], s["angle2"])


@pytest.mark.gpu
@pytest.mark.parametrize("seed,ratio,nodes", [(1, 0.95, 4), (2, 0.7, 1), (3, 0.99, 16)])
def test_search_by_bow_dense_competition(seed, ratio, nodes):
    """Few distinct descriptors in big nodes: ties everywhere, every feature's best candidates are taken by earlier ones — the replay
    kernel's conflict rounds, exhausted candidate caches and rescans all fire."""
    import orb_slam_cuda_b200 as X
    rng = np.random.default_rng(seed)
    n1, n2 = 700, 650
    base = rng.integers(0, 256, (12, 32), dtype=np.uint8)

    def table(n):
        d = base[rng.integers(0, 12, n)].copy()
        d[:, :2] ^= rng.integers(0, 256, (n, 2), dtype=np.uint8)  # up to 16 flipped bits: close neighbours, few exact ties
        return d

    d1, d2 = table(n1), table(n2)
    node1, node2 = rng.integers(0, nodes, n1), rng.integers(0, nodes, n2)
    fv1 = {k: np.flatnonzero(node1 == k).astype(np.uint32) for k in range(nodes)}
    fv2 = {k: np.flatnonzero(node2 == k).astype(np.uint32) for k in range(nodes)}
    s1, i1, s2, i2 = common_nodes(fv1, fv2)
    a1, a2 = rng.uniform(0, 360, n1).astype(np.float32), rng.uniform(0, 360, n2).astype(np.float32)
    v1, v2 = (rng.random(n1) < 0.9).astype(np.uint8), (rng.random(n2) < 0.9).astype(np.uint8)
    for ori in (False, True):
        ref = O.search_by_bow(s1, i1, s2, i2, v1, d1, a1, v2, d2, a2, 50, ratio, ori)
        got = X.ORBmatcher(ratio, ori).SearchByBoW(s1, i1, s2, i2, v1, d1, a1, v2, d2, a2, 50)
        assert got[0] == ref[0] and np.array_equal(got[1], ref[1]) and np.array_equal(got[2], ref[2])
    assert O.search_by_bow(s1, i1, s2, i2, v1, d1, a1, v2, d2, a2, 50, ratio, False)[0] > 100
