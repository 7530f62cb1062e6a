"""The oracle's extractor against the REFERENCE's own object code: /root/reference/src/ORBextractor.cc compiled where it lies
(oracle/ref_extractor/Makefile -> oracle/_ref/libref_extractor.so; CPU branch of operator(), do-nothing OpenVX, the five OpenCV
pixel primitives = the oracle's cv2-pinned models, a monotonic heap for quadtree nodes so that the reference's pointer
tie-breaks are defined — oracle/ref_extractor/ref_extractor.cpp).  Bit-exact: keypoint order, coordinates, size, angle,
response, octave and all 32 descriptor bytes.  Live where /root/reference exists, through tests/golden/ref_extractor.npz
(tools/gen_extractor_golden.py) everywhere."""
import os

import numpy as np
import pytest

from oracle import oracle_lib as O
from oracle.ref_extractor import ref_extractor as R
import ref_extractor_cases as RC

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "ref_extractor.npz")
live = pytest.mark.skipif(not R.available(), reason="reference sources absent (GPU box): covered by the golden fixture")


def same_keypoints(a, b):
    assert len(a) == len(b), (len(a), len(b))
    for f in ("x", "y", "size", "angle", "response", "octave", "class_id"):
        assert np.array_equal(a[f], b[f]), f


@live
@pytest.mark.parametrize("name", list(RC.CASES))
def test_oracle_equals_reference_object_code(name):
    make, kw = RC.CASES[name]
    img = make()
    h, w = img.shape
    ref = R.RefExtractor(w, h, **kw)
    orc = O.Extractor(**RC.oracle_kwargs(kw))
    rk, rd = ref.extract(img)
    ok, od = orc.extract(img)
    same_keypoints(rk, ok)
    assert np.array_equal(rd, od)
    for l in range(orc.nlevels):  # ComputePyramid, src/ORBextractor.cc:1837-1862
        assert np.array_equal(ref.level(l), orc.level_padded(l)[19:-19, 19:-19]), l
    if name == "flat":
        assert len(rk) == 0
    else:
        assert len(rk) > 0.5 * min(kw["nfeatures"], 500)


@live
def test_constructor_tables():
    ref, orc = R.RefExtractor(1241, 376, 2000), O.Extractor(2000)
    tr, to = ref.tables(), orc.tables()
    assert np.array_equal(tr["quota"], to["quota"])  # src/ORBextractor.cc:521-533
    assert np.array_equal(tr["scale"], to["scale"])
    assert np.array_equal(tr["umax"], to["umax"])  # :540-555
    # the 512 pattern points as the reference COMPILES them (bit_pattern_31_ holds "VX_FAILURE - 2" at element 96)
    assert np.array_equal(tr["pattern"], O.pattern_values())  # default variant = PATTERN_REFERENCE
    assert tr["pattern"][96] == -3


@live
def test_second_frame_same_object():
    """operator() twice on one extractor (mvImagePyramid is reused): the second frame does not see the first"""
    ref, orc = R.RefExtractor(640, 480, 1000), O.Extractor(1000)
    for k in (3, 4, 3):
        img = RC.synth.frame("tum", k)
        rk, rd = ref.extract(img)
        ok, od = orc.extract(img)
        same_keypoints(rk, ok)
        assert np.array_equal(rd, od)


@live
@pytest.mark.parametrize("seed,n,N", [(1, 3000, 400), (2, 500, 450), (3, 5000, 1000), (4, 40, 100), (5, 2000, 37)])
def test_distribute_octtree_ties(seed, n, N):
    """DistributeOctTree (src/ORBextractor.cc:889-1120) on clouds with few distinct responses and clustered positions: the
    equal-size-node tie-break and the stop-in-the-middle-of-a-pass rule decide the result"""
    r = np.random.default_rng(seed)
    keys = np.zeros(n, O.KP_DTYPE)
    W, H = 1209, 344
    cx, cy = r.integers(0, W, 12), r.integers(0, H, 12)
    which = r.integers(0, 12, n)
    keys["x"] = np.clip(cx[which] + r.integers(-40, 41, n), 0, W - 1)
    keys["y"] = np.clip(cy[which] + r.integers(-40, 41, n), 0, H - 1)
    keys["response"] = r.integers(1, 6, n) * 10
    keys["size"] = 7
    ref = R.RefExtractor(1241, 376, 2000)
    a = ref.distribute(keys, 0, W, 0, H, N)
    b = O.distribute(keys, 0, W, 0, H, N)
    assert len(a) == len(b) > 0
    assert np.array_equal(a[:, 0], b["x"]) and np.array_equal(a[:, 1], b["y"]) and np.array_equal(a[:, 2], b["response"])


@pytest.mark.parametrize("name", RC.GOLDEN)
def test_oracle_against_reference_golden(name):
    g = np.load(GOLDEN)
    make, kw = RC.CASES[name]
    img = make()
    assert int(g[name + "/img_sum"]) == int(img.astype(np.int64).sum())  # the fixture was made from this image
    ok, od = O.Extractor(**RC.oracle_kwargs(kw)).extract(img)
    rk = g[name + "/kps"]
    assert len(rk) == len(ok)
    for i, f in enumerate(("x", "y", "size", "angle", "response")):
        assert np.array_equal(rk[:, i], ok[f]), f
    assert np.array_equal(rk[:, 5].astype(np.int32), ok["octave"])
    assert np.array_equal(g[name + "/desc"], od)
