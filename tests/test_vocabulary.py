"""DBoW2 vocabulary tree descent — Frame::ComputeBoW (reference src/Frame.cc:394-401) -> TemplatedVocabulary::transform
(Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1127-1194, :1218-1261) and loadFromTextFile (:1338-1424); SURVEY.md §8f rank 2.
ORBvoc.txt itself is not in the reference checkout (.MISSING_LARGE_BLOBS), so the trees here are synthetic: same format, same
k = 10, shallower, with early leaves, stopped words (weight 0) and ties.  CPU: known answers for the oracle.  GPU: the CUDA
descent and the text loader through the C ABI against the oracle, bit-exact (word / node ids, weights and BowVector doubles)."""
import numpy as np
import pytest

from oracle import oracle_lib as O


def make_vocabulary(seed, k=10, L=4, p_early=0.08, p_stop=0.05, scoring=0, weighting=0, shuffle=False):
    """A tree numbered the way HKmeansStep creates it (siblings get consecutive ids, then each child is expanded,
    TemplatedVocabulary.h:642-760); shuffle=True renumbers so that siblings are scattered (any parent-before-child order loads)."""
    rng = np.random.default_rng(seed)
    parent, leaf, desc, weight = [-1], [0], [np.zeros(32, np.uint8)], [0.0]

    def expand(pid, level):
        kids = []
        for _ in range(k if rng.random() > 0.1 else int(rng.integers(2, k + 1))):
            if level == 1:
                d = rng.integers(0, 256, 32, dtype=np.uint8)
            else:
                flips = np.zeros(256, np.uint8)
                flips[rng.choice(256, size=int(96 / level), replace=False)] = 1
                d = desc[pid] ^ np.packbits(flips)
            is_leaf = level == L or (level > 1 and rng.random() < p_early)
            parent.append(pid)
            leaf.append(int(is_leaf))
            desc.append(d)
            weight.append(0.0 if (is_leaf and rng.random() < p_stop) else float(rng.uniform(0.1, 9.0)) if is_leaf else 0.0)
            kids.append((len(parent) - 1, is_leaf))
        for nid, is_leaf in kids:
            if not is_leaf:
                expand(nid, level + 1)

    expand(0, 1)
    parent, leaf, desc, weight = np.array(parent, np.int32), np.array(leaf, np.uint8), np.stack(desc), np.array(weight, np.float64)
    if shuffle:
        # new id = rank of (depth, random): parents still precede children, siblings no longer adjacent
        depth = np.zeros(parent.size, np.int64)
        for i in range(1, parent.size):
            depth[i] = depth[parent[i]] + 1
        order = np.lexsort((rng.random(parent.size), depth))
        new_id = np.empty(parent.size, np.int64)
        new_id[order] = np.arange(parent.size)
        parent = np.where(parent[order] >= 0, new_id[np.maximum(parent[order], 0)], -1).astype(np.int32)
        leaf, desc, weight = leaf[order], desc[order], weight[order]
    return O.Vocabulary(k, L, scoring, weighting, parent, leaf, desc, weight)


def features(seed, voc, n=1500):
    """Descriptors near the leaves (so descents go deep) plus pure noise plus exact copies of node descriptors (ties)."""
    rng = np.random.default_rng(seed + 1000)
    src = voc.desc[rng.integers(1, voc.desc.shape[0], n)]
    noise = rng.integers(0, 256, (n, 32), dtype=np.uint8) & rng.integers(0, 256, (n, 32), dtype=np.uint8) & \
        rng.integers(0, 256, (n, 32), dtype=np.uint8)
    f = src ^ noise
    f[::7] = rng.integers(0, 256, (len(f[::7]), 32), dtype=np.uint8)
    f[::11] = src[::11]
    return f


def test_oracle_kat_descent():
    # root -> {1, 2}; 2 = leaf (word 0: word ids follow node order), an early leaf at level 1; 1 -> {3 (word 1), 4 (word 2)}
    z, ff = [0] * 32, [255] * 32
    d3 = [0] * 32
    d4 = [15] + [0] * 31
    voc = O.Vocabulary(2, 2, 0, 0, [-1, 0, 0, 1, 1], [0, 0, 1, 1, 1], [z, z, ff, d3, d4], [0, 0, 2.5, 1.25, 0.75])
    feat = np.array([z, d4, ff, [3] + [0] * 31], np.uint8)
    word, w, node = voc.transform_features(feat, levelsup=0)      # node level 2
    assert list(word) == [1, 2, 0, 1] and list(w) == [1.25, 0.75, 2.5, 1.25]
    assert list(node) == [3, 4, 0xFFFFFFFF, 3]                    # feature 2 stops at level 1: nid unwritten in the reference
    assert list(voc.transform_features(feat, levelsup=1)[2]) == [1, 1, 2, 1]
    assert list(voc.transform_features(feat, levelsup=2)[2]) == [0, 0, 0, 0]
    # a tie goes to the first child: [3] is 2 bits from d3 and 2 bits from d4
    assert voc.transform_features(np.array([[3] + [0] * 31], np.uint8), 0)[0][0] == 1
    bow, fv = voc.transform(feat, levelsup=0)
    total = 1.25 + 0.75 + 2.5 + 1.25
    assert bow == {0: 2.5 / total, 1: (1.25 + 1.25) / total, 2: 0.75 / total}
    assert fv == {3: [0, 3], 4: [1, 2]}                           # feature 2 inherits the previous feature's node (stale nid)


def test_oracle_text_round_trip(tmp_path):
    voc = make_vocabulary(5, k=4, L=3)
    for trailing in (True, False):
        path = tmp_path / f"voc{int(trailing)}.txt"
        voc.to_text(path, trailing_newline=trailing)
        back = O.Vocabulary.from_text(path)
        # blank lines (the remainder after the last newline) make no node: the reference reads pid uninitialised there
        assert np.array_equal(back.parent, voc.parent) and np.array_equal(back.desc, voc.desc)
        assert np.array_equal(back.weight, voc.weight) and np.array_equal(back.is_leaf, voc.is_leaf)
        assert (back.k, back.L, back.scoring, back.weighting) == (4, 3, 0, 0)


GOLDEN_CASES = [(31, 6, 3, 0, 0), (32, 4, 4, 1, 1), (33, 5, 3, 5, 0), (34, 6, 2, 0, 2), (35, 3, 5, 5, 3)]


@pytest.mark.parametrize("case", range(len(GOLDEN_CASES)))
def test_oracle_against_reference_dbow2_golden(case, tmp_path):
    """The pin: tests/golden/dbow2_transform.npz holds the outputs of the REFERENCE's own DBoW2 (compiled from
    /root/reference/Thirdparty/DBoW2 by oracle/ref_dbow2/Makefile, run by tools/gen_dbow2_golden.py) on text vocabularies;
    the oracle loads the same text and must reproduce word ids, weights, node ids, BowVector doubles and FeatureVector lists."""
    import os
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "dbow2_transform.npz"))
    seed, k, L, scoring, weighting = GOLDEN_CASES[case]
    path = tmp_path / "voc.txt"
    path.write_bytes(g[f"c{case}_text"].tobytes())
    voc = O.Vocabulary.from_text(path)
    assert (voc.k, voc.L, voc.scoring, voc.weighting) == (k, L, scoring, weighting)
    assert int(voc.is_leaf[1:].sum()) == int(g[f"c{case}_size"][0])
    f = g[f"c{case}_feat"]
    for lu in range(0, L + 2):
        word, w, node = voc.transform_features(f, lu)
        assert np.array_equal(word, g[f"c{case}_lu{lu}_word"]) and np.array_equal(node, g[f"c{case}_lu{lu}_node"])
        assert np.array_equal(w.view(np.uint64), g[f"c{case}_lu{lu}_weight"].view(np.uint64))
    assert (g[f"c{case}_lu0_node"] == 0xFFFFFFFF).any() or L < 3   # the golden run met early leaves (nid left unwritten)
    bow, fv = voc.transform(f, 2)
    assert np.array_equal(np.array(list(bow.keys()), np.uint32), g[f"c{case}_bow_word"])
    assert np.array_equal(np.array(list(bow.values()), np.float64).view(np.uint64), g[f"c{case}_bow_value"].view(np.uint64))
    assert np.array_equal(np.array(list(fv.keys()), np.uint32), g[f"c{case}_fv_node"])
    assert np.array_equal(np.array([len(v) for v in fv.values()], np.int32), g[f"c{case}_fv_count"])
    assert np.array_equal(np.array([i for v in fv.values() for i in v], np.uint32), g[f"c{case}_fv_index"])


def test_oracle_against_compiled_reference_dbow2(tmp_path):
    """Where the compiled reference is present (this container; oracle/_ref travels to the GPU box), fresh random cases."""
    from oracle.ref_dbow2 import ref_dbow2 as R
    if R.lib() is None:
        pytest.skip("oracle/_ref/libref_dbow2.so not built (needs /root/reference)")
    for seed, (k, L, scoring, weighting) in enumerate([(10, 4, 0, 0), (7, 3, 1, 0), (2, 7, 0, 1), (9, 3, 5, 2)], start=50):
        voc = make_vocabulary(seed, k, L, p_early=0.1, p_stop=0.08, scoring=scoring, weighting=weighting, shuffle=seed % 2 == 1)
        path = tmp_path / f"voc{seed}.txt"
        voc.to_text(path, trailing_newline=False)   # with the newline the reference reads an uninitialised parent id
        ref = R.RefVocabulary(path)
        assert ref.size() == int(voc.is_leaf[1:].sum())
        f = features(seed, voc, 600)
        for lu in (0, 1, 2, 4):
            a, b = voc.transform_features(f, lu), ref.transform_features(f, lu)
            assert np.array_equal(a[0], b[0]) and np.array_equal(a[1].view(np.uint64), b[1].view(np.uint64)) and np.array_equal(a[2], b[2])
        bow, fv = voc.transform(f, 2)
        rbow, rfv = ref.transform(f, 2)
        assert list(bow.items()) == list(rbow.items()) and fv == rfv


@pytest.mark.parametrize("scoring,weighting", [(0, 0), (1, 1), (5, 0), (0, 2), (5, 3)])
def test_oracle_bow_vector_normalisation(scoring, weighting):
    voc = make_vocabulary(2, k=5, L=3, scoring=scoring, weighting=weighting)
    bow, fv = voc.transform(features(2, voc, 400), 2)
    vals = np.array(list(bow.values()))
    assert len(bow) > 20 and sum(len(v) for v in fv.values()) <= 400
    if scoring == 0:
        assert abs(vals.sum() - 1.0) < 1e-12
    elif scoring == 1:
        assert abs((vals ** 2).sum() - 1.0) < 1e-12
    elif weighting == 0:
        assert vals.sum() > 0 and abs(vals.sum() - 1.0) > 1e-6     # DOT_PRODUCT: divided by the number of words, not normalised


def voc_gpu(X, voc):
    return X.ORBVocabulary(voc.k, voc.L, voc.scoring, voc.weighting, voc.parent, voc.is_leaf, voc.desc, voc.weight)


@pytest.mark.gpu
@pytest.mark.parametrize("seed,k,L,shuffle", [(1, 10, 4, False), (2, 10, 3, True), (3, 3, 6, False), (4, 20, 2, True)])
def test_vocabulary_transform_gpu(seed, k, L, shuffle):
    import orb_slam_cuda_b200 as X
    voc = make_vocabulary(seed, k, L, shuffle=shuffle)
    v = voc_gpu(X, voc)
    info = v.info()
    assert info["n_nodes"] == voc.parent.size and info["n_words"] == int(voc.is_leaf[1:].sum()) and info["k"] == k and info["L"] == L
    f = features(seed, voc)
    for levelsup in (0, 1, 2, 4, 10):
        r_word, r_w, r_node = voc.transform_features(f, levelsup)
        word, w, node = v.transform_features(f, levelsup)
        assert np.array_equal(word, r_word) and np.array_equal(node, r_node)
        assert np.array_equal(w.view(np.uint64), r_w.view(np.uint64))
    assert L < 3 or (voc.transform_features(f, 0)[2] == 0xFFFFFFFF).any()   # early leaves are exercised
    assert (voc.transform_features(f, 4)[1] == 0.0).any()           # and so are stopped words


@pytest.mark.gpu
@pytest.mark.parametrize("scoring,weighting", [(0, 0), (1, 1), (5, 0), (0, 2), (5, 3)])
def test_compute_bow_gpu(scoring, weighting):
    """Frame::ComputeBoW on descriptors of a real extraction: BowVector doubles and FeatureVector lists equal the oracle's."""
    import orb_slam_cuda_b200 as X
    from orb_slam_cuda_b200 import synth
    voc = make_vocabulary(11, 10, 4, scoring=scoring, weighting=weighting)
    _, d = O.Extractor(1000, 1.2, 8, 20, 7).extract(synth.frame("tum", 3))
    # real ORB descriptors are far from a random tree's centroids; fold half of them onto node descriptors so the descents spread
    d = d.copy()
    rng = np.random.default_rng(0)
    half = d.shape[0] // 2
    d[:half] = voc.desc[rng.integers(1, voc.desc.shape[0], half)] ^ (d[:half] & d[half:2 * half] & 0x11)
    ref_bow, ref_fv = voc.transform(d, 2)
    bow, fv = voc_gpu(X, voc).transform(d, 2)
    assert len(ref_bow) > 50 and list(bow.keys()) == list(ref_bow.keys())
    assert all(np.float64(bow[k]).view(np.uint64) == np.float64(ref_bow[k]).view(np.uint64) for k in bow)
    assert fv == ref_fv


@pytest.mark.gpu
def test_vocabulary_text_loader_gpu(tmp_path):
    import orb_slam_cuda_b200 as X
    voc = make_vocabulary(6, 10, 3)
    path = tmp_path / "ORBvoc.txt"
    voc.to_text(path)
    ref = O.Vocabulary.from_text(path)
    v = X.ORBVocabulary.loadFromTextFile(path)
    info = v.info()
    assert info["n_nodes"] == ref.parent.size == voc.parent.size and info["n_words"] == int(ref.is_leaf[1:].sum())
    assert (info["k"], info["L"], info["scoring"], info["weighting"]) == (10, 3, 0, 0)
    f = features(6, voc)
    r = ref.transform_features(f, 1)
    g = v.transform_features(f, 1)
    assert all(np.array_equal(a, b) for a, b in zip(g, r))
    # without the trailing newline, and with blank lines in between: the same tree
    path2 = tmp_path / "voc2.txt"
    path2.write_text(path.read_text().rstrip("\n").replace("\n", "\n\n", 3))
    v2 = X.ORBVocabulary.loadFromTextFile(path2)
    assert v2.info() == info and all(np.array_equal(a, b) for a, b in zip(v2.transform_features(f, 1), r))
    # a line that starts but does not parse
    path3 = tmp_path / "voc3.txt"
    path3.write_text(path.read_text() + "7 1 2 3\n")
    with pytest.raises(X.OrbxError):
        X.ORBVocabulary.loadFromTextFile(path3)
    # not a vocabulary / missing file / malformed tree
    bad = tmp_path / "bad.txt"
    bad.write_text("hello world\n")
    for p in (bad, tmp_path / "missing.txt"):
        with pytest.raises(X.OrbxError):
            X.ORBVocabulary.loadFromTextFile(p)
    with pytest.raises(X.OrbxError):
        X.ORBVocabulary(2, 1, 0, 0, [-1, 2, 0], [0, 1, 1], np.zeros((3, 32), np.uint8), [0, 1, 1])   # parent after child
    with pytest.raises(X.OrbxError):
        X.ORBVocabulary(2, 1, 0, 0, [-1, 1], [0, 1], np.zeros((2, 32), np.uint8), [0, 1])            # root without children
