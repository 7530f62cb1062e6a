"""CPU suite, part 2: the oracle against cv2 4.13 LIVE (skipped where cv2 is not importable), as a second opinion:
the five OpenCV primitives on many shapes, and a Python transliteration of the reference's control flow
(src/ORBextractor.cc CPU path) built on the cv2 primitives, compared stage by stage with the C++ oracle.
Also pins the device's restatement of glibc sinf / cosf (csrc/k_describe.cu) against the host libm."""
import math

import numpy as np
import pytest

from oracle import oracle_lib as O
from orb_slam_cuda_b200 import synth

cv2 = pytest.importorskip("cv2")
cv2.setNumThreads(1)
f32 = np.float32
H = float.fromhex


@pytest.mark.parametrize("shape,dsize", [((376, 1241), (1034, 313)), ((126, 416), (346, 105)), ((480, 640), (533, 400)),
                                         ((97, 211), (176, 81)), ((60, 80), (40, 30)), ((31, 45), (44, 30))])
def test_resize_live(shape, dsize):
    img = synth.textured_frame(shape[0] * 7 + shape[1], *shape)
    assert np.array_equal(O.resize_linear(img, *dsize), cv2.resize(img, dsize, interpolation=cv2.INTER_LINEAR))


@pytest.mark.parametrize("shape", [(376, 1241), (105, 346), (40, 57), (9, 11)])
def test_blur_and_border_live(shape):
    img = synth.textured_frame(shape[0] + 3 * shape[1], *shape)
    assert np.array_equal(O.gaussian7(img), cv2.GaussianBlur(img, (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101))
    if min(shape) > 19:
        assert np.array_equal(O.border_reflect101(img, 19), cv2.copyMakeBorder(img, 19, 19, 19, 19, cv2.BORDER_REFLECT_101))


def _fast_cv2(img, th):
    det = cv2.FastFeatureDetector_create(threshold=th, nonmaxSuppression=True, type=cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
    return [(k.pt[0], k.pt[1], k.response) for k in det.detect(img, None)]


@pytest.mark.parametrize("th", [20, 7])
def test_fast_live(th):
    img = synth.textured_frame(99, 240, 320)
    got = O.fast9(img, th, True)
    want = _fast_cv2(img, th)
    assert len(want) > 100
    assert [(float(k["x"]), float(k["y"]), float(k["response"])) for k in got] == want
    # small windows like the reference's per-cell calls (36 x 36 and clipped ones)
    for (y, x, h, w) in [(10, 20, 36, 36), (200, 280, 36, 37), (0, 0, 7, 7), (50, 60, 8, 30)]:
        sub = np.ascontiguousarray(img[y:y + h, x:x + w])
        got = O.fast9(sub, th, True)
        assert [(float(k["x"]), float(k["y"]), float(k["response"])) for k in got] == _fast_cv2(sub, th)


def test_fast_atan2_live():
    rng = np.random.default_rng(3)
    y = rng.integers(-2_700_000, 2_700_000, 20000).astype(f32)
    x = rng.integers(-2_700_000, 2_700_000, 20000).astype(f32)
    want = cv2.phase(x, y, angleInDegrees=True).ravel()
    got = O.fast_atan2(y, x)
    scalar = np.array([cv2.fastAtan2(float(a), float(b)) for a, b in zip(y[:3000], x[:3000])], f32)
    assert np.array_equal(got[:3000].view(np.uint32), scalar.view(np.uint32))
    assert np.abs(got - want).max() < 1e-3  # the vectorised cv::phase may differ from the scalar path in the last bits


# ---- Python transliteration of the reference's control flow on cv2 primitives -------------------------------------
def _cv_round(v):
    return int(np.rint(v))  # round half to even


class PyRefExtractor:
    """src/ORBextractor.cc CPU path, restated in Python over cv2 (ctor :496-557, ComputePyramid :1837-1862,
    ComputeKeyPointsOctTree :1122-1340 CPU branch, DistributeOctTree :889-1120, operator() :1701-1809)."""

    def __init__(self, nfeatures, scaleFactor, nlevels, ini, mn):
        self.nf, self.nl, self.ini, self.mn = nfeatures, nlevels, ini, mn
        sfd = float(f32(scaleFactor))
        self.scale = [f32(1.0)]
        for _ in range(1, nlevels):
            self.scale.append(f32(float(self.scale[-1]) * sfd))
        self.inv = [f32(1.0) / s for s in self.scale]
        factor = f32(1.0 / sfd)
        nd = f32(nfeatures) * (f32(1) - factor) / (f32(1) - f32(math.pow(float(factor), nlevels)))
        self.quota, tot = [], 0
        for _ in range(nlevels - 1):
            self.quota.append(_cv_round(nd))
            tot += self.quota[-1]
            nd = f32(nd * factor)
        self.quota.append(max(nfeatures - tot, 0))

    def pyramid(self, image):
        E = 19
        self.pad, self.lvl = [], []
        for l in range(self.nl):
            w = _cv_round(f32(image.shape[1]) * self.inv[l])
            h = _cv_round(f32(image.shape[0]) * self.inv[l])
            if l == 0:
                lvl = image
            else:
                lvl = cv2.resize(self.lvl[l - 1], (w, h), interpolation=cv2.INTER_LINEAR)
            self.lvl.append(lvl)
            self.pad.append(cv2.copyMakeBorder(lvl, E, E, E, E, cv2.BORDER_REFLECT_101))

    def candidates(self, l):
        E, W = 19, f32(30)
        pad = self.pad[l]
        lw, lh = self.lvl[l].shape[1], self.lvl[l].shape[0]
        minBX = minBY = E - 3
        maxBX, maxBY = lw - E + 3, lh - E + 3
        width, height = f32(maxBX - minBX), f32(maxBY - minBY)
        nCols, nRows = int(width / W), int(height / W)
        if nCols < 1 or nRows < 1:
            return [], (minBX, maxBX, minBY, maxBY)
        wCell, hCell = int(math.ceil(width / nCols)), int(math.ceil(height / nRows))
        out = []
        for i in range(nRows):
            iniY = minBY + i * hCell
            maxY = iniY + hCell + 6
            if iniY >= maxBY - 3:
                continue
            maxY = min(maxY, maxBY)
            for j in range(nCols):
                iniX = minBX + j * wCell
                maxX = iniX + wCell + 6
                if iniX >= maxBX - 6:
                    continue
                maxX = min(maxX, maxBX)
                sub = np.ascontiguousarray(pad[E + iniY:E + maxY, E + iniX:E + maxX])
                kps = _fast_cv2(sub, self.ini)
                if not kps:
                    kps = _fast_cv2(sub, self.mn)
                out += [(x + j * wCell, y + i * hCell, r) for (x, y, r) in kps]
        return out, (minBX, maxBX, minBY, maxBY)


def py_distribute(keys, minX, maxX, minY, maxY, N):
    """DistributeOctTree (:889-1120) with Python lists standing in for std::list; ties of the final-phase sort are
    broken by creation order, as in the oracle (the reference ties on heap addresses)."""
    nIni = int(math.floor(abs(float(f32(maxX - minX) / f32(maxY - minY))) + 0.5))
    hX = f32(maxX - minX) / f32(nIni)
    seq = [0]

    def node(x0, y0, x1, y1):
        seq[0] += 1
        return {"x0": x0, "y0": y0, "x1": x1, "y1": y1, "keys": [], "seq": seq[0]}

    L = [node(int(hX * f32(i)), 0, int(hX * f32(i + 1)), maxY - minY) for i in range(nIni)]
    for k in keys:
        L[int(f32(k[0]) / hX)]["keys"].append(k)
    L = [n for n in L if n["keys"]]

    def divide(p):
        hx = int(math.ceil(f32(p["x1"] - p["x0"]) / 2))
        hy = int(math.ceil(f32(p["y1"] - p["y0"]) / 2))
        mx, my = p["x0"] + hx, p["y0"] + hy
        ch = [node(p["x0"], p["y0"], mx, my), node(mx, p["y0"], p["x1"], my), node(p["x0"], my, mx, p["y1"]),
              node(mx, my, p["x1"], p["y1"])]
        for k in p["keys"]:
            if k[0] < mx:
                ch[0 if k[1] < my else 2]["keys"].append(k)
            else:
                ch[1 if k[1] < my else 3]["keys"].append(k)
        return [c for c in ch if c["keys"]]

    finish = False
    while not finish:
        prev = len(L)
        i, nToExpand, expandable = 0, 0, []
        while i < len(L):
            if len(L) >= N:
                finish = True
                break
            n = L[i]
            if len(n["keys"]) == 1:
                i += 1
                continue
            kids = divide(n)
            del L[i]
            for c in kids:
                L.insert(0, c)
                i += 1
                if len(c["keys"]) > 1:
                    nToExpand += 1
                    expandable.append(c)
        if len(L) >= N or len(L) == prev:
            finish = True
        elif len(L) + 3 * nToExpand > N:
            while not finish:
                prev = len(L)
                todo = sorted(expandable, key=lambda c: (len(c["keys"]), c["seq"]))
                expandable = []
                for n in reversed(todo):
                    kids = divide(n)
                    for c in kids:
                        L.insert(0, c)
                        if len(c["keys"]) > 1:
                            expandable.append(c)
                    L.remove(n)
                    if len(L) >= N:
                        break
                if len(L) >= N or len(L) == prev:
                    finish = True
    out = []
    for n in L:
        best = n["keys"][0]
        for k in n["keys"][1:]:
            if k[2] > best[2]:
                best = k
        out.append(best)
    return out


@pytest.mark.parametrize("shape,nf", [((240, 320), 500), ((200, 331), 800)])
def test_pipeline_transliteration(shape, nf):
    img = synth.textured_frame(555 + shape[1], *shape)
    ox = O.Extractor(nf, 1.2, 8, 20, 7)
    okps, odesc = ox.extract(img)
    py = PyRefExtractor(nf, 1.2, 8, 20, 7)
    assert py.quota == list(ox.tables()["quota"])
    py.pyramid(img)
    total = 0
    for l in range(8):
        assert np.array_equal(py.pad[l], ox.level_padded(l)), f"pyramid level {l}"
        cand, (minBX, maxBX, minBY, maxBY) = py.candidates(l)
        oc = ox.level_candidates(l)
        assert cand == [(float(k["x"]), float(k["y"]), float(k["response"])) for k in oc], f"FAST candidates level {l}"
        kept = py_distribute(cand, minBX, maxBX, minBY, maxBY, py.quota[l]) if cand else []
        ok = ox.level_keypoints(l)
        assert [(x + minBX, y + minBY, r) for (x, y, r) in kept] == \
            [(float(k["x"]), float(k["y"]), float(k["response"])) for k in ok], f"quadtree level {l}"
        total += len(kept)
        if len(kept) == 0:
            continue
        blur = cv2.GaussianBlur(py.lvl[l], (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101)
        assert np.array_equal(blur, ox.level_blurred(l)), f"blur level {l}"
        # IC_Angle (:164-191) on a sample of keypoints, moments in int, cv2.fastAtan2
        umax = list(ox.tables()["umax"])
        pad = py.pad[l].astype(np.int64)
        for k in ok[:: max(1, len(ok) // 25)]:
            cx, cy = int(k["x"]) + 19, int(k["y"]) + 19
            m10 = m01 = 0
            for v in range(-15, 16):
                d = umax[abs(v)]
                row = pad[cy + v, cx - d:cx + d + 1]
                m10 += int((np.arange(-d, d + 1) * row).sum())
                m01 += v * int(row.sum())
            assert f32(cv2.fastAtan2(float(m01), float(m10))) == k["angle"], (l, cx, cy)
    assert total == len(okps) and total > nf // 2
    # scale-back and level-major order (:1794-1803)
    o = 0
    for l in range(8):
        for k in ox.level_keypoints(l):
            s = py.scale[l]
            want = (f32(k["x"]) * s, f32(k["y"]) * s) if l else (k["x"], k["y"])
            assert (okps[o]["x"], okps[o]["y"], okps[o]["octave"]) == (want[0], want[1], l)
            o += 1


def test_descriptor_transliteration():
    """computeOrbDescriptor (:195-233) in numpy float32 with libm cosf / sinf taken from the oracle's binding."""
    img = synth.textured_frame(777, 200, 260)
    ox = O.Extractor(400, 1.2, 8, 20, 7)
    okps, odesc = ox.extract(img)
    inc = []
    import os
    for line in open(os.path.join(os.path.dirname(os.path.abspath(synth.__file__)), "csrc", "orb_pattern_31.inc")):
        if not line.startswith("//"):
            inc += [int(t) for t in line.replace(",", " ").split()]
    inc[96] = -3  # the reference as it compiles: VX_FAILURE - 2 (src/ORBextractor.cc:261-262); the oracle's default variant
    pat = np.array(inc, np.float32).reshape(512, 2)
    o = 0
    for l in range(8):
        lk = ox.level_keypoints(l)
        blur = ox.level_blurred(l)
        for k in lk:
            if o % 9 == 0:
                ang = f32(k["angle"]) * f32(math.pi / 180.0)
                c, s = O.sincosf(np.array([ang], f32))
                a, b = c[0], s[0]
                rows = np.rint(pat[:, 0] * b + pat[:, 1] * a).astype(np.int64)  # float32 products, separate rounding
                cols = np.rint(pat[:, 0] * a - pat[:, 1] * b).astype(np.int64)
                vals = blur[int(k["y"]) + rows, int(k["x"]) + cols].reshape(256, 2)
                bits = (vals[:, 0] < vals[:, 1]).astype(np.uint8)
                assert np.array_equal(np.packbits(bits, bitorder="little"), odesc[o]), (l, o)
            o += 1
    assert o == len(okps)


# ---- the device restatement of glibc sinf / cosf (csrc/k_describe.cu glibc_sincosf) -------------------------------
def _glibc_sincosf_model(y):
    """numpy float64 transliteration of the kernel's code path (every operation individually rounded)."""
    y = y.astype(np.float32)
    x = y.astype(np.float64)
    S1, S2, S3 = -H("0x1.555545995a603p-3"), H("0x1.1107605230bc4p-7"), -H("0x1.994eb3774cf24p-13")
    C0, C1, C2, C3, C4 = H("0x1p0"), -H("0x1.ffffffd0c621cp-2"), H("0x1.55553e1068f19p-5"), -H("0x1.6c087e89a359dp-10"), H("0x1.99343027bf8c3p-16")

    def poly(xs, x2, n, neg):
        sg = np.where(neg, -1.0, 1.0)
        x3 = xs * x2
        s1 = S2 + x2 * S3
        x7 = x3 * x2
        s = xs + x3 * S1
        sinv = s + x7 * s1
        x4 = x2 * x2
        c2 = sg * C3 + x2 * (sg * C4)
        c1 = sg * C0 + x2 * (sg * C1)
        x6 = x4 * x2
        c = c1 + x4 * (sg * C2)
        cosv = c + x6 * c2
        return np.where((n & 1) == 0, sinv, cosv).astype(np.float32)

    top = (y.view(np.uint32) >> 20) & 0x7FF
    small, tiny = top < 0x3F4, top < 0x398
    r = x * H("0x1.45F306DC9C883p+23")
    n = (np.trunc(r).astype(np.int64) + 0x800000) >> 24
    xr = x + -(n.astype(np.float64) * H("0x1.921FB54442D18p0"))
    sgn = np.where(((n & 3) == 1) | ((n & 3) == 2), -1.0, 1.0)
    neg = (n & 2) != 0
    xs = xr * sgn
    sn_big, cs_big = poly(xs, xr * xr, n, neg), poly(xs, xr * xr, n ^ 1, neg)
    zero, no = np.zeros_like(n), np.zeros(n.shape, bool)
    sn_small, cs_small = poly(x, x * x, zero, no), poly(x, x * x, zero + 1, no)
    sn = np.where(tiny, y, np.where(small, sn_small, sn_big)).astype(np.float32)
    cs = np.where(tiny, np.float32(1.0), np.where(small, cs_small, cs_big)).astype(np.float32)
    return sn, cs


def test_sincosf_restatement_matches_libm():
    rng = np.random.default_rng(11)
    deg = np.concatenate([rng.uniform(0, 360, 2_000_000).astype(np.float32), np.arange(0, 360, 0.25, dtype=np.float32),
                          np.float32([0, 1e-5, 1e-3, 44.990456, 90, 180, 270, 359.99997])])
    ang = deg * np.float32(math.pi / 180.0)
    c, s = O.sincosf(ang)
    sn, cs = _glibc_sincosf_model(ang)
    assert np.array_equal(sn.view(np.uint32), s.view(np.uint32)), f"{(sn != s).sum()} sines differ"
    assert np.array_equal(cs.view(np.uint32), c.view(np.uint32)), f"{(cs != c).sum()} cosines differ"
