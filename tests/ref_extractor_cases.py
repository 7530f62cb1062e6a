"""Seeded cases shared by tests/test_oracle_ref_extractor.py and tools/gen_extractor_golden.py: the reference's own
ORBextractor object code (oracle/_ref/libref_extractor.so) against the oracle's restatement, whole frames.
Each case: name -> (image, constructor arguments)."""
import numpy as np

from orb_slam_cuda_b200 import synth


def _noise(seed, h, w):
    return np.random.default_rng(seed).integers(0, 256, (h, w), dtype=np.uint8)


def _blocks(seed, h, w, cell=24):
    """flat blocks with sharp corners: many equal FAST responses, so many equal-size quadtree nodes (the tie-break rule)"""
    r = np.random.default_rng(seed)
    g = r.integers(0, 4, ((h + cell - 1) // cell, (w + cell - 1) // cell), dtype=np.uint8) * 60 + 20
    return np.ascontiguousarray(np.kron(g, np.ones((cell, cell), np.uint8))[:h, :w])


def _dim(seed, h, w):
    """low contrast: most cells fall through to the minThFAST pass (src/ORBextractor.cc:1281-1285)"""
    return (synth.textured_frame(seed, h, w).astype(np.int32) // 6 + 100).astype(np.uint8)


def _strided(seed, h, w):
    """a view with a row stride larger than its width (cv::Mat step != cols)"""
    big = synth.textured_frame(seed, h, w + 40)
    return big[:, 8:8 + w]


CASES = {
    # the three configurations of BASELINE.json (KITTI00-02.yaml, TUM1.yaml, EuRoC.yaml), two frames each
    "kitti_0": (lambda: synth.frame("kitti", 0), dict(nfeatures=2000)),
    "kitti_5": (lambda: synth.frame("kitti", 5), dict(nfeatures=2000)),
    "tum_1": (lambda: synth.frame("tum", 1), dict(nfeatures=1000)),
    "tum_7": (lambda: synth.frame("tum", 7), dict(nfeatures=1000)),
    "euroc_2": (lambda: synth.frame("euroc", 2), dict(nfeatures=1200)),
    "euroc_9": (lambda: synth.frame("euroc", 9), dict(nfeatures=1200)),
    # the monocular initialisation extractor (src/Tracking.cc: 2 * nFeatures)
    "kitti_init": (lambda: synth.frame("kitti", 2), dict(nfeatures=4000)),
    # other pyramids and thresholds
    "levels4_sf15": (lambda: synth.textured_frame(11, 300, 400), dict(nfeatures=800, scale_factor=1.5, nlevels=4)),
    "levels1": (lambda: synth.textured_frame(12, 240, 320), dict(nfeatures=300, nlevels=1)),
    "thresholds": (lambda: synth.textured_frame(13, 360, 480), dict(nfeatures=700, ini_th=35, min_th=12)),
    "few": (lambda: synth.textured_frame(14, 360, 480), dict(nfeatures=60)),
    # content that stresses one branch
    "noise": (lambda: _noise(15, 300, 420), dict(nfeatures=1500)),
    "blocks": (lambda: _blocks(16, 360, 480), dict(nfeatures=1000)),
    "dim": (lambda: _dim(17, 360, 480), dict(nfeatures=1000)),
    "flat": (lambda: np.full((300, 400), 128, np.uint8), dict(nfeatures=500)),
    "half_flat": (lambda: np.hstack([np.full((376, 600), 90, np.uint8), synth.textured_frame(18, 376, 641)]), dict(nfeatures=2000)),
    "strided": (lambda: _strided(19, 300, 400), dict(nfeatures=900)),
    "odd_size": (lambda: synth.textured_frame(20, 331, 517), dict(nfeatures=1100)),
}
GOLDEN = ["kitti_0", "tum_1", "euroc_2", "levels4_sf15", "noise", "blocks", "dim", "few"]  # stored in tests/golden/ref_extractor.npz


def oracle_kwargs(kw):
    return dict(nfeatures=kw.get("nfeatures", 2000), scale_factor=kw.get("scale_factor", 1.2), nlevels=kw.get("nlevels", 8),
                ini_th=kw.get("ini_th", 20), min_th=kw.get("min_th", 7))
