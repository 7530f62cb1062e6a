"""Frame::UndistortKeyPoints (reference src/Frame.cc:403-433) — SURVEY.md §8f rank 4.  CPU: the oracle's restatement of
cv::undistortPoints against cv2.undistortPoints itself (live) and a committed fixture.  GPU: the kernel through the C ABI against the
oracle, bit-exact (float outputs compared as raw bits, tolerance 0)."""
import os

import numpy as np
import pytest

from oracle import oracle_lib as O

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "undistort.npz")
# k1 k2 p1 p2 [k3] of the reference's example cameras: Examples/Monocular/TUM1.yaml:12-16, TUM2.yaml, EuRoC.yaml (mono), plus KITTI (none)
CAMERAS = {
    "tum1": ((517.306408, 516.469215, 318.643040, 255.313989), (0.262383, -0.953104, -0.005358, 0.002628, 1.163314), (640, 480)),
    "tum2": ((520.908620, 521.007327, 325.141442, 249.701764), (0.231222, -0.784899, -0.003257, -0.000105, 0.917205), (640, 480)),
    "euroc": ((458.654, 457.296, 367.215, 248.375), (-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05), (752, 480)),
    "kitti": ((718.856, 718.856, 607.1928, 185.2157), (0.0, 0.0, 0.0, 0.0), (1241, 376)),
}


def keypoints(name, n=3000, seed=1):
    _, _, (w, h) = CAMERAS[name]
    rng = np.random.default_rng(seed)
    k = np.zeros(n, O.KP_DTYPE)
    k["x"], k["y"] = rng.uniform(0, w, n).astype(np.float32), rng.uniform(0, h, n).astype(np.float32)
    k["x"][:4], k["y"][:4] = [0, w, 0, w], [0, 0, h, h]  # the image corners (ComputeImageBounds, src/Frame.cc:435-462)
    k["size"], k["angle"], k["response"], k["octave"], k["class_id"] = 31.0, rng.uniform(0, 360, n), rng.integers(7, 255, n), rng.integers(0, 8, n), -1
    return k


@pytest.mark.parametrize("name", list(CAMERAS))
def test_oracle_against_cv2_live(name):
    cv2 = pytest.importorskip("cv2")
    K4, dist, _ = CAMERAS[name]
    k = keypoints(name)
    got = O.undistort_keypoints(k, K4, dist)
    if dist[0] == 0:
        assert np.array_equal(got.view(np.uint8), k.view(np.uint8))
        return
    K = np.array([[K4[0], 0, K4[2]], [0, K4[1], K4[3]], [0, 0, 1]], np.float32)
    pts = np.stack([k["x"], k["y"]], 1).reshape(-1, 1, 2).astype(np.float32)
    want = cv2.undistortPoints(pts, K, np.array(dist, np.float32), None, K).reshape(-1, 2)
    assert np.array_equal(got["x"].view(np.uint32), want[:, 0].view(np.uint32))
    assert np.array_equal(got["y"].view(np.uint32), want[:, 1].view(np.uint32))
    for f in ("size", "angle", "response", "octave", "class_id"):
        assert np.array_equal(got[f], k[f])
    assert np.abs(got["x"] - k["x"]).max() > 1.0  # the model does move the points


@pytest.mark.parametrize("name", [n for n in CAMERAS if CAMERAS[n][1][0] != 0])
def test_oracle_against_cv2_golden(name):
    g = np.load(GOLD)
    K4, dist, _ = CAMERAS[name]
    got = O.undistort_keypoints(keypoints(name), K4, dist)
    assert np.array_equal(np.stack([got["x"], got["y"]], 1).view(np.uint32), g[name])


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(CAMERAS))
def test_undistort_gpu(name):
    import orb_slam_cuda_b200 as X
    K4, dist, _ = CAMERAS[name]
    k = keypoints(name, n=2500, seed=7)
    got = X.UndistortKeyPoints(k, K4, dist)
    want = O.undistort_keypoints(k, K4, dist)
    assert np.array_equal(got.view(np.uint8), want.view(np.uint8))
    assert X.UndistortKeyPoints(k[:0], K4, dist).size == 0
