"""The oracle against the REFERENCE's own object code: /root/reference/src/ORBmatcher.cc and src/Frame.cc compiled where they lie
(oracle/ref_matcher/Makefile -> oracle/_ref/libref_matcher.so) and driven through tests/ref_matcher_cases.py.  Covers
DescriptorDistance, Frame::GetFeaturesInArea + AssignFeaturesToGrid, SearchForInitialization, SearchByProjection (frame to frame,
local map, Sim3), SearchByBoW (both overloads), SearchForTriangulation, Fuse (candidate loop AND map surgery) and
Frame::ComputeStereoMatches — integer results equal, float results (mvuRight, mvDepth, vbPrevMatched) equal bit for bit.
`live` needs the library (built in the container that has /root/reference; the built .so travels to the GPU box); `golden`
runs anywhere from tests/golden/ref_matcher.npz (tools/gen_matcher_golden.py)."""
import os

import numpy as np
import pytest

import ref_matcher_cases as RC

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_matcher.npz")


def _same(a, b, name):
    assert len(a) == len(b), name
    for i, (x, y) in enumerate(zip(a, b)):
        assert x.shape == y.shape and np.array_equal(x, y), f"{name}: output {i} differs from the reference"


@pytest.mark.parametrize("name", list(RC.CASES))
def test_oracle_against_compiled_reference_live(name):
    if not RC.R.available():
        pytest.skip("oracle/_ref/libref_matcher.so is not here and /root/reference is absent: the golden test covers this box")
    oracle, reference = RC.CASES[name]()
    out = oracle()
    assert out[0].size > 0
    _same(out, reference(), name)


@pytest.mark.parametrize("name", list(RC.CASES))
def test_oracle_against_reference_golden(name):
    g = np.load(GOLDEN)
    oracle, _ = RC.CASES[name]()
    out = oracle()
    ref = [g[f"{name}/{i}"] for i in range(len(out))]
    assert f"{name}/{len(out)}" not in g.files
    _same(out, ref, name)


def test_cases_are_not_trivial():
    """Every case produces matches / operations of both kinds, so equality is not vacuous."""
    g = np.load(GOLDEN)
    assert g["sfi_tum/0"].min() > 50 and g["sbp_frame_mono/0"][0] > 300 and g["bow_kf_kf/0"][0] > 100
    assert g["triangulation/0"][0] > 100 and g["sbp_sim3/0"][0] > 100
    ops = g["fuse_stereo/1"]
    assert (ops[:, 0] == 2).sum() > 20 and (ops[:, 0] == 0).sum() > 20 and g["fuse_stereo/0"][0] > 100
    ur = g["stereo_euroc/0"].view(np.float32)
    assert (ur >= 0).sum() > 300
