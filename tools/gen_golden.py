#!/usr/bin/env python3
"""Generates tests/golden/*.npz — the fixtures that pin the CPU oracle (development-time tool; cv2 4.13 needed).

The reference ships no tests or golden vectors, and its sources do not compile here, so the
oracle's pixel arithmetic is pinned against the OpenCV primitives the reference calls (resize :1852, copyMakeBorder
:1854-1860, FAST :1278-1284, GaussianBlur :1747, fastAtan2 :190 of src/ORBextractor.cc), evaluated by the cv2 4.13
wheel of this image on small seeded inputs.  The outputs are committed so the tests do not need cv2, and
tests/test_oracle_cv2.py re-checks them live where cv2 is importable.

  cv2_primitives.npz : inputs + cv2 outputs of the five primitives
  oracle_frame.npz   : a 200x152 synthetic frame + the oracle's own keypoints / descriptors (regression vector, made by
                       the oracle AFTER it passed the cv2 checks; guards the control flow against accidental edits)
"""
import os
import sys

import cv2
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from orb_slam_cuda_b200 import synth  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")
os.makedirs(OUT, exist_ok=True)


def fast_cv2(img, th):
    det = cv2.FastFeatureDetector_create(threshold=th, nonmaxSuppression=True, type=cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
    kps = det.detect(img, None)
    return np.array([(k.pt[0], k.pt[1], k.response) for k in kps], dtype=np.float32).reshape(-1, 3)


def main():
    cv2.setNumThreads(1)
    img = synth.textured_frame(424242, 96, 131)
    out = {"img": img}
    # resize: the KITTI L0 -> L1 ratio on a small image, an odd ratio, and an exact 2x (INTER_AREA shortcut of cv::resize)
    for name, (dw, dh) in {"a": (109, 80), "b": (77, 41), "c": (130, 95)}.items():
        out[f"resize_{name}"] = cv2.resize(img, (dw, dh), interpolation=cv2.INTER_LINEAR)
    img2 = synth.textured_frame(424243, 96, 132)
    out["img2"] = img2
    out["resize_half"] = cv2.resize(img2, (66, 48), interpolation=cv2.INTER_LINEAR)
    out["border19"] = cv2.copyMakeBorder(img, 19, 19, 19, 19, cv2.BORDER_REFLECT_101)
    out["blur"] = cv2.GaussianBlur(img, (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101)
    tiny = img[:9, :11].copy()
    out["tiny"] = tiny
    out["blur_tiny"] = cv2.GaussianBlur(tiny, (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101)
    out["fast20"] = fast_cv2(img, 20)
    out["fast7"] = fast_cv2(img, 7)
    rng = np.random.default_rng(7)
    y = rng.integers(-2_700_000, 2_700_000, 4000).astype(np.float32)
    x = rng.integers(-2_700_000, 2_700_000, 4000).astype(np.float32)
    y[:8] = [0, 0, -5, 5, 1, -1, 0, 3]
    x[:8] = [0, -5, 0, 0, 1, -1, 7, 3]
    out["atan_y"], out["atan_x"] = y, x
    out["atan"] = np.array([cv2.fastAtan2(float(a), float(b)) for a, b in zip(y, x)], dtype=np.float32)
    np.savez_compressed(os.path.join(OUT, "cv2_primitives.npz"), **out)

    from oracle import oracle_lib as O
    frame = synth.textured_frame(20261104, 152, 200)
    ox = O.Extractor(300, 1.2, 8, 20, 7)
    kps, desc = ox.extract(frame)
    np.savez_compressed(os.path.join(OUT, "oracle_frame.npz"), frame=frame, kps=kps, desc=desc,
                        params=np.array([300, 8, 20, 7], np.int32))
    # cv2.undistortPoints on the keypoint sets of tests/test_undistort.py (Frame::UndistortKeyPoints, src/Frame.cc:403-433)
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
    import test_undistort as TU
    und = {}
    for name, (K4, dist, _) in TU.CAMERAS.items():
        if dist[0] == 0:
            continue
        k = TU.keypoints(name)
        K = np.array([[K4[0], 0, K4[2]], [0, K4[1], K4[3]], [0, 0, 1]], np.float32)
        pts = np.stack([k["x"], k["y"]], 1).reshape(-1, 1, 2).astype(np.float32)
        und[name] = cv2.undistortPoints(pts, K, np.array(dist, np.float32), None, K).reshape(-1, 2).view(np.uint32)
    np.savez_compressed(os.path.join(OUT, "undistort.npz"), **und)
    print("wrote", OUT, "fast20", len(out["fast20"]), "fast7", len(out["fast7"]), "oracle kps", len(kps))


if __name__ == "__main__":
    main()
