"""cvtColor to gray ahead of the extractor — reference src/Tracking.cc:221-246, :261-274, :294-307.
CPU: the oracle's fixed-point restatement against cv2 4.13 (all four conversion codes).  GPU: the kernel through the C ABI against
the oracle, bit-exact, for both coefficient sets, odd widths and 3 / 4 channels."""
import numpy as np
import pytest

from oracle import oracle_lib as O


def test_oracle_against_cv2():
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(1)
    for shape in [(37, 53), (480, 640), (1, 1), (5, 4099)]:
        for ch, codes in [(3, (cv2.COLOR_RGB2GRAY, cv2.COLOR_BGR2GRAY)), (4, (cv2.COLOR_RGBA2GRAY, cv2.COLOR_BGRA2GRAY))]:
            img = rng.integers(0, 256, shape + (ch,), dtype=np.uint8)
            img[0, 0, :3] = 255
            for bgr, code in enumerate(codes):
                assert np.array_equal(O.cvt_color_gray(img, bool(bgr)), cv2.cvtColor(img, code))
    # the OpenCV 2.4 / 3.x constants differ by at most one grey level
    img = rng.integers(0, 256, (200, 300, 3), dtype=np.uint8)
    a, b = O.cvt_color_gray(img).astype(int), O.cvt_color_gray(img, cv2_coeffs=True).astype(int)
    assert np.abs(a - b).max() == 1 and (a != b).mean() < 0.01


@pytest.mark.gpu
@pytest.mark.parametrize("shape,ch", [((376, 1241), 3), ((480, 640), 4), ((33, 7), 3), ((2, 4098), 4), ((1, 1), 3)])
def test_cvt_color_gray_gpu(shape, ch):
    import orb_slam_cuda_b200 as X
    rng = np.random.default_rng(2)
    img = rng.integers(0, 256, shape + (ch,), dtype=np.uint8)
    for bgr in (False, True):
        for old in (False, True):
            assert np.array_equal(X.cvtColorGray(img, bgr, old), O.cvt_color_gray(img, bgr, old))
    with pytest.raises(X.OrbxError):
        X.cvtColorGray(np.zeros((4, 4, 2), np.uint8))
    # gray -> extractor: the converted frame gives the same keypoints as the oracle on the oracle's conversion
    if shape == (480, 640):
        k, d = X.ORBextractor(500, 1.2, 8, 20, 7)(X.cvtColorGray(img, True))
        ok, od = O.Extractor(500, 1.2, 8, 20, 7).extract(O.cvt_color_gray(img, True))
        assert np.array_equal(k.view(np.uint8), ok.view(np.uint8)) and np.array_equal(d, od)
